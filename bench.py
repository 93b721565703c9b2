#!/usr/bin/env python
"""bench.py -- headline measurement of the block-sparse product path (BASELINE.json metric:
SpGEMM useful GFLOP/s; chain milliseconds as a secondary figure).

    python bench.py --gpus N --steps K --warmup W            our CUDA path
    python bench.py --impl reference --gpus N --steps K ...  the reference's own CPU path

A step is one pass of the hot path (symbolic + numeric) over one synthetic workload:
BASELINE.json configs[1], a single A*B with n=65536, k=8, 1 % block density, uint64 entries.
At N > 1 the SAME product is sharded over the N GPUs by contiguous block-row panels balanced by
tile products (bsp_panel_bounds; B replicated, no data-path collective):
total work is fixed ("strong") and value = products of the whole job / max-over-ranks time.
Useful ops = 2*k^3*nprod (one multiply + one add per inner step), SURVEY.md section 8d.
Every figure carries the content checksum of its result and is compared with the CPU oracle's
checksum of the same full-size workload (tests/golden/full_size.json, made by
tests/golden/make_full_size.py); a mismatch ends the run with a non-zero exit code.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200")
sys.path.insert(0, PKG)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (n, k, density, value kind)
    "cfg2": (65536, 8, 0.01, "u64"),
    "cfg1": (1024, 4, 0.05, "u64"),
    "mid": (16384, 8, 0.01, "u64"),
    "k16": (65536, 16, 0.02, "u64"),
    "k32": (65536, 32, 0.02, "u64"),   # the tile size of the reference's own published numbers (report.pdf)
}


def ncu_traffic(workload, path_id):
    """DRAM bytes of one numeric-kernel launch from the latest committed `ncu --set full` capture
    (profiles/ncu_traffic.json names the capture, its date and the commit it was taken at)."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            for e in json.load(f)["captures"]:
                if e["workload"] == workload and e["numeric_path"] == path_id:
                    return e["dram_bytes_read"] + e["dram_bytes_write"], "%s (%s, commit %s)" % (e["file"], e["date"], e["commit"])
    except (OSError, KeyError, ValueError):
        pass
    return None, None


def int_mac_peak():
    """u64 multiply-accumulate ceiling of the numeric kernel's inner loop (tools/u64mac_peak.cu run on a B200)."""
    try:
        best = 0.0
        with open(os.path.join(ROOT, "profiles", "r01_u64mac_peak.jsonl")) as f:
            for ln in f:
                e = json.loads(ln)
                if e.get("variant") == "flagged":
                    best = max(best, float(e["u64_mac_per_s"]))
        return best or None
    except (OSError, ValueError):
        return None


def load_golden():
    try:
        with open(os.path.join(ROOT, "tests", "golden", "full_size.json")) as f:
            return json.load(f)
    except (OSError, ValueError):
        return {}


class ParityError(SystemExit):
    pass


def check_golden(what, got_hex, want_hex):
    """A bench figure whose result differs from the oracle's is not a figure: stop with exit code 3."""
    if want_hex is None:
        return "no golden for this configuration"
    if got_hex != want_hex:
        sys.stderr.write("bench.py: PARITY FAILURE in %s: checksum %s, oracle %s\n" % (what, got_hex, want_hex))
        raise ParityError(3)
    return "equal to the oracle's (tests/golden/full_size.json)"
METRIC = "spgemm_useful_gflops"
UNIT = "GFLOP/s"


def workload_name(n, k, d):
    return "single A*B, n=%d, k=%d, %.3g%% block density, uint64 entries (BASELINE.json configs[1] shape)" % (n, k, d * 100)


def algorithmic_bytes(nnzb_a, nnzb_b, nnzb_c, G, k):
    """Compulsory CSR-of-blocks traffic of the numeric pass."""
    per = lambda nnzb: nnzb * k * k * 8 + 4 * nnzb + 4 * (G + 1)  # noqa: E731
    return per(nnzb_a) + per(nnzb_b) + per(nnzb_c)


def symbolic_bytes(nnzb_a, nprod, nnzb_c, G):
    """A block columns, the B block columns of every row touched, C block columns, row pointers."""
    return 4 * nnzb_a + 4 * nprod + 4 * nnzb_c + 12 * (G + 1)


def pair_list_bytes(nprod, nnzb_c):
    """What this design writes on top of 8d: 8 B per tile product and one 8 B offset per output tile."""
    return 8 * nprod + 8 * (nnzb_c + 1)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.rows.append([x.strip() for x in ln.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); smax.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for nm, val in zip(names, r[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def gpu_local_affinity(torch, local):
    """Restrict this process to the CPUs NVML reports as local to GPU `local` (its NUMA node), so that
    the pinned host buffers of the end-to-end leg are allocated next to the GPU's PCIe root; a
    remote node costs 20-30 % of the transfer rate.  Returns (previous mask, note) -- (None, why)
    when nothing was changed."""
    try:
        import pynvml
        pynvml.nvmlInit()
        pr = torch.cuda.get_device_properties(local)
        bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64 + 16)
        near = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        old = os.sched_getaffinity(0)
        new = near & old
        if not new or new == old:
            return None, "unchanged (%d allowed CPUs, %d of them GPU-local)" % (len(old), len(new))
        os.sched_setaffinity(0, new)
        return old, "%d GPU-local CPUs of %d allowed" % (len(new), len(old))
    except Exception as e:  # NVML missing, cpuset restrictions, ...
        return None, "unchanged (%s)" % type(e).__name__


# --------------------------------------------------------------------------------------------
# the reference arm / cpu_baseline: the reference's own translation unit built host-only
# --------------------------------------------------------------------------------------------

def reference_sample(n, k, d, seed_a, seed_b, rows_frac=8, inner_frac=8):
    """A bounded sample of the workload for the CPU arm: block rows [0, G/rows_frac) of A and
    inner block indices [0, G/inner_frac) (columns of A / rows of B), all columns of B."""
    import bsp_gen
    G = n // k
    R, J = max(1, G // rows_frac), max(1, G // inner_frac)
    _, _, _, abr, abc, av = bsp_gen.random_matrix(n, n, k, d, seed_a, "u64")
    _, _, _, bbr, bbc, bv = bsp_gen.random_matrix(n, n, k, d, seed_b, "u64")
    ka = (abr < R) & (abc < J)
    kb = bbr < J
    a = (R * k, J * k, k, abr[ka], abc[ka], av[ka])
    b = (J * k, n, k, bbr[kb], bbc[kb], bv[kb])
    nprod = int(np.bincount(b[3], minlength=J)[a[4]].sum())
    desc = ("block rows [0,%d) of A x inner block indices [0,%d) x all %d block columns of B: %d x %d tiles, %d tile products"
            % (R, J, G, a[3].size, b[3].size, nprod))
    return a, b, nprod, desc


def run_reference_once(folder, variant, threads):
    """One timed run of the reference program; returns hot-path seconds (sum of its own phase
    timers inside helper, cu:162-273) and its end-to-end 'time taken' (cu:679)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    r = orc.run_ref(folder, ranks=1, variant=variant, threads=threads, keep_output=False)
    if r["returncode"] != 0:
        raise RuntimeError("reference run failed: " + r["stderr"][-500:])
    hot = sum(r["phases"].values())
    return hot, (r["time_taken"][0] if r["time_taken"] else None), r["phases"]


def cpu_reference_measure(n, k, d, steps, warmup, use_all_threads=True):
    """Times the reference's CPU path on a bounded sample. Returns dict for the JSON line."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    cores = os.cpu_count() or 1
    a, b, nprod, desc = reference_sample(n, k, d, 1000, 1001)
    flop = 2.0 * k ** 3 * nprod
    if orc.have_ref("ref_host_timed_omp"):
        variant = "ref_host_timed_omp" if use_all_threads else "ref_host_timed"
        threads = cores if use_all_threads else 1
        with tempfile.TemporaryDirectory(prefix="bsp_cpu_") as folder:
            orc.write_folder(folder, [orc.OMat.from_coo(*a), orc.OMat.from_coo(*b)], k)
            times, e2e = [], []
            for i in range(warmup + steps):
                hot, tt, _ = run_reference_once(folder, variant, threads)
                if i >= warmup:
                    times.append(hot); e2e.append(tt)
        sec = sum(times) / len(times)
        return dict(value=flop / sec / 1e9, unit=UNIT, cores=threads, kind="reference",
                    sample=desc + "; hot path = sum of the reference's own phase timers inside helper; "
                    "kernel loop emulated on %d OpenMP threads, the rest single-threaded as shipped" % threads,
                    seconds_hot_path=sec, seconds_end_to_end=(sum(e2e) / len(e2e)) if e2e and e2e[0] else None,
                    nprod=nprod)
    # the reference host build did not travel: time the C restatement instead
    A, B = orc.OMat.from_coo(*a), orc.OMat.from_coo(*b)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        orc.multiply(A, B, threads=cores)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    sec = sum(times) / len(times)
    return dict(value=flop / sec / 1e9, unit=UNIT, cores=cores, kind="port", sample=desc, seconds_hot_path=sec, nprod=nprod)


REF_CHAIN = dict(n_mats=64, n=1024, k=4, half_band=4, density=0.5)  # the chain of CHAIN at 1/16 of its block rows


def reference_chain_measure(ranks):
    """The reference program (its own main: MPI split over `ranks` ranks through the fork+pipe shim,
    helper2 per rank, rank-0 merge) on a REDUCED copy of the chain workload -- same generator, same
    band, 64 matrices, n=1024 instead of 16384: the full chain's last product alone is 1.8e8 tile
    products, hours of single-threaded host packing (cu:189-218).  Returns hot-path and end-to-end ms."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import bsp_gen
    import orc
    c = REF_CHAIN
    G = c["n"] // c["k"]
    mats = []
    for i in range(c["n_mats"]):
        br, bc = bsp_gen.banded_pattern(G, c["half_band"], c["density"], 5000 + i)
        v = bsp_gen.values(5000 + i, br.size * c["k"] * c["k"], "small").reshape(-1, c["k"], c["k"])
        mats.append(orc.OMat.from_coo(c["n"], c["n"], c["k"], br, bc, v))
    variant = "ref_host_timed" if orc.have_ref("ref_host_timed") else None
    if variant is None:
        t0 = time.perf_counter()
        orc.chain(mats, ranks=ranks)
        sec = time.perf_counter() - t0
        return {"ms": sec * 1e3, "kind": "port", "ranks": ranks, "workload": "reduced chain %r" % (c,)}
    with tempfile.TemporaryDirectory(prefix="bsp_cpu_chain_") as folder:
        orc.write_folder(folder, mats, c["k"])
        r = orc.run_ref(folder, ranks=ranks, variant=variant, keep_output=False)
    if r["returncode"] != 0:
        raise RuntimeError("reference chain run failed: " + r["stderr"][-300:])
    return {"ms": (max(r["time_taken"]) if r["time_taken"] else float("nan")) * 1e3, "hot_path_ms_rank0": sum(r["phases"].values()) * 1e3,
            "kind": "reference", "ranks": ranks, "products": len(r["multiplying"]),
            "workload": "reduced chain %r, values in [1,1000]; ms = the reference's own 'time taken' line (cu:679, includes its text I/O)" % (c,)}


def reference_cfg1_measure():
    """BASELINE.json configs[0] in full (n=1024, k=4, 5 %): the one configuration the reference runs as it is."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import bsp_gen
    import orc
    n, k, d, _ = WORKLOADS["cfg1"]
    a = bsp_gen.random_matrix(n, n, k, d, 1000, "u64")
    b = bsp_gen.random_matrix(n, n, k, d, 1001, "u64")
    nprod = int(np.bincount(b[3], minlength=n // k)[a[4]].sum())
    if not orc.have_ref("ref_host_timed"):
        t0 = time.perf_counter()
        orc.multiply(orc.OMat.from_coo(*a), orc.OMat.from_coo(*b))
        sec = time.perf_counter() - t0
        return {"gflops": 2.0 * k ** 3 * nprod / sec / 1e9, "hot_path_ms": sec * 1e3, "kind": "port", "nprod": nprod}
    with tempfile.TemporaryDirectory(prefix="bsp_cpu_cfg1_") as folder:
        orc.write_folder(folder, [orc.OMat.from_coo(*a), orc.OMat.from_coo(*b)], k)
        hot, tt, _ = run_reference_once(folder, "ref_host_timed", 1)
    return {"gflops": 2.0 * k ** 3 * nprod / hot / 1e9, "hot_path_ms": hot * 1e3, "end_to_end_ms": tt * 1e3 if tt else None,
            "kind": "reference", "nprod": nprod, "workload": "configs[0] in full: n=%d, k=%d, %g%% density, single rank as shipped" % (n, k, d * 100)}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n, k, d, _ = WORKLOADS[args.workload]
    steps = max(1, min(args.steps, 5))
    warm = min(args.warmup, 1)
    cb = cpu_reference_measure(n, k, d, steps, warm)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": warm, "ms_per_step": cb["seconds_hot_path"] * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": workload_name(n, k, d), "sample": cb["sample"],
                       "note": "the reference cannot shard one product (SURVEY.md 8e): at every N this line is its single-rank "
                               "rate on the sample; its N-rank figure is the chain below"},
            "cpu_baseline": {kk: cb[kk] for kk in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    for key, fn in (("chain", lambda: reference_chain_measure(max(1, args.gpus))), ("cfg1", reference_cfg1_measure)):
        try:
            line[key] = fn()
        except Exception as ex:  # reported baselines, never load-bearing
            line[key] = {"unavailable": str(ex)[:200]}
    print(json.dumps(line))
    return 0


# --------------------------------------------------------------------------------------------
# secondary figure: chain milliseconds (BASELINE.json configs[2])
# --------------------------------------------------------------------------------------------
CHAIN = dict(n_mats=64, n=16384, k=4, half_band=4, density=0.5, kind="u64")


def chain_matrix(i):
    """Matrix i of the chain workload: banded block pattern (|brow - bcol| <= half_band, each
    slot kept with probability `density`), so fill-in stays bounded: the product of m factors has
    half-band m * half_band.  BASELINE.json gives no density for the chain; this is our choice."""
    import bsp_gen
    c = CHAIN
    G = c["n"] // c["k"]
    br, bc = bsp_gen.banded_pattern(G, c["half_band"], c["density"], 5000 + i)
    v = bsp_gen.values(5000 + i, br.size * c["k"] * c["k"], c["kind"]).reshape(-1, c["k"], c["k"])
    return c["n"], c["n"], c["k"], br, bc, v


def make_comm(torch, dist, ctx, rank, world):
    """One NCCL communicator of the C ABI per rank (bsp_comm_create); the rendezvous token travels over
    torch.distributed, which the bench has anyway for its barriers."""
    import bspgemm
    tok = torch.zeros(bspgemm.COMM_ID_BYTES, dtype=torch.uint8, device="cuda")
    if rank == 0:
        tok.copy_(torch.frombuffer(bytearray(bspgemm.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(tok, 0)
    torch.cuda.synchronize()
    return bspgemm.Comm(ctx, bytes(tok.cpu().numpy().tobytes()), world, rank)


def summed_checksum(torch, dist, world, mat, rows, cols, k):
    """Checksum of a matrix spread over the ranks as row panels: tile sums and tile counts add up."""
    import bspgemm
    body = mat.checksum_body() if mat is not None else 0
    nnzb = mat.nnzb if mat is not None else 0
    v = torch.tensor([body - (1 << 64) if body >= (1 << 63) else body, nnzb], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(v, op=dist.ReduceOp.SUM)  # int64 addition wraps: the sum modulo 2^64
    body_all, nnzb_all = int(v[0].item()) & ((1 << 64) - 1), int(v[1].item())
    return "%016x" % bspgemm.checksum_finish(rows, cols, k, 8, nnzb_all, body_all), nnzb_all


def run_chain(torch, dist, ctx, comm, rank, world, steps, warmup, golden):
    """Chain of 64 matrices split into contiguous sub-chains over the ranks exactly as the
    reference splits it over MPI ranks (cu:438-456), local pairwise tree per rank, then the
    step-doubling merge of the partial products through the C ABI (bsp_chain_distributed: NCCL
    send/recv, every product sharded over its subtree).  Strong scaling: the chain is fixed.
    Returns dict for the JSON line (rank 0) or None."""
    import bsp_dist
    import bspgemm
    parts = bsp_dist.split_chain(CHAIN["n_mats"], world)
    mine = parts[rank] if rank < len(parts) else None
    mats = [ctx.matrix_from_coo(*chain_matrix(i)) for i in range(mine[0], mine[1] + 1)] if mine else []
    nprod = 0

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one():
        nonlocal nprod
        st = bspgemm.Stats()
        if world == 1:
            out = ctx.helper2(mats, stats=st)
        else:
            out = comm.chain(mats, stats=st, log_base=mine[0] if mine else 0)
        nprod = st.nprod
        return out

    res = None
    for _ in range(warmup):
        res = one()
    sync()
    sent0 = comm.bytes_sent if comm else 0
    stream = torch.cuda.current_stream()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    evs[0].record(stream)
    for i in range(steps):
        res = one()
        evs[i + 1].record(stream)
    sync()
    per_step = [evs[i].elapsed_time(evs[i + 1]) for i in range(steps)]
    sent = ((comm.bytes_sent - sent0) // steps) if comm else 0
    t = torch.tensor([evs[0].elapsed_time(evs[steps]) / steps], dtype=torch.float64, device="cuda")
    w = torch.tensor([float(nprod)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(w, op=dist.ReduceOp.SUM)
    for m in mats:
        m.free()
    if rank != 0:
        return None
    k = CHAIN["k"]
    cs = "%016x" % res.checksum()
    want = golden.get("chain", {}).get("ranks", {}).get(str(len(parts)), {})
    verdict = check_golden("chain (%d ranks)" % len(parts), cs, want.get("checksum"))
    out = {"ms": t.item(), "unit": "ms", "n_gpus": world, "scaling": "strong", "steps": steps, "warmup": warmup,
           "per_step_ms_rank0": [round(x, 3) for x in per_step],
           "workload": "chain of %d matrices, n=%d, k=%d, banded pattern (half-band %d blocks, %.0f%% filled), uint64; "
                       "sub-chains of %d per GPU, merge by bsp_chain_distributed (NCCL send/recv, every product row-panel "
                       "sharded over its subtree, panels stay distributed)" % (CHAIN["n_mats"], CHAIN["n"], k, CHAIN["half_band"],
                                                                              CHAIN["density"] * 100, CHAIN["n_mats"] // max(len(parts), 1)),
           "products": CHAIN["n_mats"] - 1, "nprod_total": int(w.item()), "result_nnzb": int(res.nnzb),
           "checksum": cs, "checksum_vs_oracle": verdict,
           "useful_gflops": 2.0 * k ** 3 * w.item() / (t.item() * 1e-3) / 1e9,
           "rank0_bytes_sent_per_step": int(sent)}
    res.free()
    return out


# --------------------------------------------------------------------------------------------
# secondary figure: power-law pattern, row-panel sharded single product (BASELINE.json configs[3])
# --------------------------------------------------------------------------------------------

def run_rmat(torch, dist, ctx, rank, world, steps, golden):
    """R-MAT block pattern (scale 15 = 32768 block rows, edge factor 16, a/b/c/d = .57/.19/.19/.05),
    n=262144, k=8.  Every rank holds A and B (same seeds), multiplies ONE block-row panel of A by B;
    panels are contiguous and balanced by tile products per block row (bsp_panel_bounds, on the
    device).  No collective on the data path; the product stays distributed by panels.  Strong
    scaling: the product is fixed."""
    import bsp_gen
    import bspgemm
    scale, k = 15, 8
    G = 1 << scale
    mats = []
    for seed in (11, 12):
        br, bc = bsp_gen.rmat_pattern(scale, 16, seed)
        v = bsp_gen.values(seed, br.size * k * k, "u64").reshape(-1, k, k)
        mats.append((G * k, G * k, k, br, bc, v))
    A, B = ctx.matrix_from_coo(*mats[0]), ctx.matrix_from_coo(*mats[1])
    del mats
    bounds = ctx.panel_bounds(A, B, world)
    st = bspgemm.Stats()
    ms, sym, C = [], [], None
    for i in range(steps + 2):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        if C is not None:
            C.free()
        C = ctx.helper(A, B, stats=st, panel=(bounds[rank], bounds[rank + 1]) if world > 1 else None)
        if i >= 2:
            ms.append(st.ms_total); sym.append(st.ms_symbolic)
    cs, nnzb_all = summed_checksum(torch, dist, world, C, G * k, G * k, k)
    C.free()
    mine = torch.tensor([sum(ms) / len(ms), float(st.nprod), sum(sym) / len(sym)], dtype=torch.float64, device="cuda")
    allv = [torch.zeros_like(mine) for _ in range(world)]
    if world > 1:
        dist.all_gather(allv, mine)
    else:
        allv = [mine]
    A.free(); B.free()
    if rank != 0:
        return None
    t = [x[0].item() for x in allv]
    nprod = sum(x[1].item() for x in allv)
    verdict = check_golden("R-MAT product", cs, golden.get("rmat", {}).get("checksum"))
    return {"ms": max(t), "unit": "ms", "n_gpus": world, "scaling": "strong",
            "workload": "R-MAT block pattern scale %d (n=%d, k=%d), A*B row-panel sharded x%d, panels balanced by tile products"
                        % (scale, G * k, k, world),
            "per_gpu_ms": [round(x, 3) for x in t], "per_gpu_symbolic_ms": [round(x[2].item(), 3) for x in allv],
            "max_over_mean": max(t) / (sum(t) / len(t)),
            "nprod_total": int(nprod), "nnzb_c_total": int(nnzb_all), "checksum": cs, "checksum_vs_oracle": verdict,
            "useful_gflops": 2.0 * k ** 3 * nprod / (max(t) * 1e-3) / 1e9}


# --------------------------------------------------------------------------------------------
# secondary figure: float32 tiles, tensor cores vs SIMT (BASELINE.json configs[4])
# --------------------------------------------------------------------------------------------

def run_float(torch, ctx, k, n, density, steps):
    """float32 tiles of edge k, block density `density`, values U(-1,1).  The reference has no
    float path; the checker is a float64 evaluation of sampled output tiles.  Reports both numeric
    kernels (tcgen05 TF32 and fp32 SIMT) and their worst error relative to sum|a||b|."""
    import bsp_gen
    import bspgemm
    G = n // k
    mats = []
    for seed in (21, 22):
        br, bc = bsp_gen.bernoulli_pattern(G, G, density, seed)
        u = bsp_gen.splitmix64(seed ^ 0xABCDEF, br.size * k * k).astype(np.float64) / 2.0 ** 64
        mats.append((n, n, k, br, bc, (2.0 * u - 1.0).astype(np.float32).reshape(-1, k, k)))
    A, B = ctx.matrix_from_coo_f32(*mats[0]), ctx.matrix_from_coo_f32(*mats[1])
    out = {"workload": "float32 tiles k=%d, n=%d, %.3g%% block density, U(-1,1) entries" % (k, n, density * 100)}
    a, b = mats
    a_key = a[3].astype(np.int64) * G + a[4]
    b_key = b[3].astype(np.int64) * G + b[4]
    for path in ("tensor", "simt"):
        ctx.set_f32_path(path)
        st = bspgemm.Stats()
        ms = []
        for i in range(steps + 2):
            C = ctx.helper(A, B, stats=st)
            if i >= 2:
                ms.append(st.ms_numeric)
            if i < steps + 1:
                C.free()
        br, bc, v = C.to_coo()
        rng = np.random.default_rng(5)
        worst = 0.0
        for t in rng.integers(0, br.size, 24):
            i, c = int(br[t]), int(bc[t])
            js = a[4][np.searchsorted(a[3], i):np.searchsorted(a[3], i + 1)]
            ref = np.zeros((k, k)); bound = np.zeros((k, k))
            for j in js:
                q = np.searchsorted(b_key, int(j) * G + c)
                if q < b_key.size and b_key[q] == int(j) * G + c:
                    At = a[5][np.searchsorted(a_key, i * G + int(j))].astype(np.float64)
                    Bt = b[5][q].astype(np.float64)
                    ref += At @ Bt; bound += np.abs(At) @ np.abs(Bt)
            worst = max(worst, float((np.abs(v[t] - ref) / np.maximum(bound, 1e-30)).max()))
        C.free()
        t_ms = sum(ms) / len(ms)
        # compulsory bytes of the numeric pass: every operand tile once, every output tile once, the pair lists once
        comp = (a[3].size + b[3].size + int(st.nnzb_c)) * k * k * 4 + 8 * int(st.nprod) + 8 * (int(st.nnzb_c) + 1)
        peak, peak_src = measured_peak()
        out[path] = {"numeric_ms": t_ms, "numeric_path": st.numeric_path, "gflops": 2.0 * k ** 3 * st.nprod / (t_ms * 1e-3) / 1e9,
                     "max_err_over_sum_abs": worst,
                     "roofline": {"bound": "hbm", "achieved": comp / (t_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                                  "frac": comp / (t_ms * 1e-3) / 1e9 / peak, "algorithmic_bytes": comp, "peak_source": peak_src}}
        out["nprod"], out["nnzb_c"], out["symbolic_ms"] = st.nprod, st.nnzb_c, st.ms_symbolic
    ctx.set_f32_path("auto")
    out["tolerance"] = "tensor (TF32 inputs, fp32 accumulate) <= 2e-3, simt (fp32 FMA) <= 2e-6, relative to sum_j |a_j||b_j|"
    A.free(); B.free()
    return out


# --------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 = min(steps, 3)")
    ap.add_argument("--no-chain", action="store_true")
    ap.add_argument("--rmat", action="store_true", help="(kept for old command lines: the R-MAT figure always runs)")
    ap.add_argument("--no-rmat", action="store_true")
    ap.add_argument("--no-float", action="store_true")
    ap.add_argument("--chain-steps", type=int, default=5, help="timed chain steps (at least 5, after 3 warm-ups)")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    # stdout carries exactly one JSON line: libraries that print there (NCCL's version banner)
    # are sent to stderr for the duration of the run
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    import bsp_gen
    import bspgemm

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    warmup = max(args.warmup, 3)
    n, k, d, kind = WORKLOADS[args.workload]
    G = n // k
    golden = load_golden()

    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        ctx = bspgemm.Context(local, stream=stream.cuda_stream)
        comm = make_comm(torch, dist, ctx, rank, world) if world > 1 else None
        a = bsp_gen.random_matrix(n, n, k, d, 1000, kind)  # A and B of the ONE product all ranks share
        b = bsp_gen.random_matrix(n, n, k, d, 1001, kind)
        A, B = ctx.matrix_from_coo(*a), ctx.matrix_from_coo(*b)
        # this rank's block-row panel (balanced by tile products on the device; whole product at N = 1)
        bounds = ctx.panel_bounds(A, B, world)
        panel = (bounds[rank], bounds[rank + 1]) if world > 1 else None

        def barrier():
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()

        st = bspgemm.Stats()
        for _ in range(warmup):
            ctx.helper(A, B, stats=st, panel=panel).free()
        barrier()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        launches0 = ctx.launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        num_ms, sym_ms = [], []
        Cres = None
        ev0.record(stream)
        for _ in range(args.steps):
            if Cres is not None:
                Cres.free()
            Cres = ctx.helper(A, B, stats=st, panel=panel)
            num_ms.append(st.ms_numeric); sym_ms.append(st.ms_symbolic)
        ev1.record(stream)
        barrier()
        clocks = sampler.stop() if rank == 0 else None
        launches = ctx.launch_count() - launches0
        ms_total = ev0.elapsed_time(ev1)
        t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
        work = torch.tensor([float(st.nprod), float(st.nnzb_c)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(work, op=dist.ReduceOp.SUM)
        ms_step = t.item() / args.steps
        nprod_all, nnzb_c_all = work[0].item(), int(work[1].item())
        value = 2.0 * k ** 3 * nprod_all / (ms_step * 1e-3) / 1e9
        # parity of what was just timed: the panels' checksums add up to the oracle's checksum of the product
        head_cs, _ = summed_checksum(torch, dist, world, Cres, n, n, k)
        Cres.free()
        head_verdict = None
        if rank == 0:
            want = golden.get(args.workload, {}).get("checksum") if args.workload == "cfg2" else None
            head_verdict = check_golden("headline product", head_cs, want)

        # end to end through the C ABI with HOST buffers: upload this rank's rows of A and all of B, multiply,
        # download the rank's part of C.  At N > 1 the ranks share one host's PCIe and memory bandwidth.
        e2e = None
        if not args.no_e2e:
            keep = (a[3] >= bounds[rank]) & (a[3] < bounds[rank + 1])
            ea = (a[0], a[1], a[2], a[3][keep], a[4][keep], a[5][keep])
            nc = int(st.nnzb_c)
            e_nprod = float(st.nprod)
            old_mask, affinity_note = gpu_local_affinity(torch, local)
            pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True)  # noqa: E731
            ha = [pin(ea[3].shape, torch.int32), pin(ea[4].shape, torch.int32), pin(ea[5].size, torch.int64)]
            hb = [pin(b[3].shape, torch.int32), pin(b[4].shape, torch.int32), pin(b[5].size, torch.int64)]
            ho = [pin((nc,), torch.int32), pin((nc,), torch.int32), pin((nc * k * k,), torch.int64)]
            ha[0].numpy()[:] = ea[3]; ha[1].numpy()[:] = ea[4]; ha[2].numpy().view(np.uint64)[:] = ea[5].reshape(-1)
            hb[0].numpy()[:] = b[3]; hb[1].numpy()[:] = b[4]; hb[2].numpy().view(np.uint64)[:] = b[5].reshape(-1)
            out = (ho[0].numpy(), ho[1].numpy(), ho[2].numpy().view(np.uint64).reshape(nc, k, k))
            av = (n, n, ha[0].numpy(), ha[1].numpy(), ha[2].numpy().view(np.uint64))
            bv = (n, n, hb[0].numpy(), hb[1].numpy(), hb[2].numpy().view(np.uint64))
            esteps = args.e2e_steps or min(args.steps, 3)
            ctx.multiply_host(k, av, bv, out=out)  # warm-up
            barrier()
            t0 = time.perf_counter()
            for _ in range(esteps):
                ctx.multiply_host(k, av, bv, out=out)
            torch.cuda.synchronize()
            dt = torch.tensor([(time.perf_counter() - t0) * 1e3], dtype=torch.float64, device="cuda")
            ew = torch.tensor([e_nprod], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX)
                dist.all_reduce(ew, op=dist.ReduceOp.SUM)
            e_ms = dt.item() / esteps
            h2d = sum(x.numel() * x.element_size() for x in ha + hb)
            d2h = sum(x.numel() * x.element_size() for x in ho)
            e2e = {"value": 2.0 * k ** 3 * ew.item() / (e_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": e_ms,
                   "steps": esteps, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "api": "bsp_multiply_host (pinned host buffers in, pinned host buffers out)",
                   "cpu_affinity": affinity_note,
                   "rows_per_rank": "all block rows" if world == 1 else
                   "each rank its block-row panel of A and all of B (bytes are rank 0's; the ranks share one host)"}
            del ha, hb, ho
            if old_mask is not None:
                os.sched_setaffinity(0, old_mask)

        A.free(); B.free()
        chain = None
        if not args.no_chain:
            chain = run_chain(torch, dist, ctx, comm, rank, world, max(5, args.chain_steps), 3, golden)

        rmat = None
        if not args.no_rmat:
            rmat = run_rmat(torch, dist, ctx, rank, world, 3, golden)

        floats = None
        if world == 1 and not args.no_float:
            floats = [run_float(torch, ctx, 32, 65536, 0.02, 3), run_float(torch, ctx, 64, 65536, 0.02, 3)]

        if rank == 0:
            peak, peak_src = measured_peak()
            # the launch of rank 0: its panel of A, all of B, its panel of C
            alg = algorithmic_bytes(st.nnzb_a if world == 1 else int(keep_count(a, bounds, rank)), st.nnzb_b, st.nnzb_c, G, k)
            k_ms = sum(num_ms) / len(num_ms)
            s_ms = sum(sym_ms) / len(sym_ms)
            ach = alg / (k_ms * 1e-3) / 1e9
            sb = symbolic_bytes(st.nnzb_a, st.nprod, st.nnzb_c, G)
            pl = pair_list_bytes(st.nprod, st.nnzb_c)
            traffic, traffic_src = ncu_traffic(args.workload, st.numeric_path) if world == 1 else (None, None)
            mac_peak = int_mac_peak()
            mac_rate = k ** 3 * st.nprod / (k_ms * 1e-3)
            line = {
                "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "u64", "data": "synthetic",
                "config": {"workload": workload_name(n, k, d),
                           "parallelism": "one product, %d block-row panel(s) balanced by tile products (B replicated, no collective)" % world,
                           "l2": "inputs (%.2f GB) and output (%.2f GB per step) exceed the 126 MB L2; no explicit flush"
                                 % ((st.nnzb_a + st.nnzb_b) * k * k * 8 / 1e9, nnzb_c_all * k * k * 8 / 1e9),
                           "nnzb_a": st.nnzb_a, "nnzb_b": st.nnzb_b, "nnzb_c": nnzb_c_all, "nprod": int(nprod_all),
                           "panel_bounds": bounds, "numeric_path": st.numeric_path, "exact_refolds": st.exact_refolds},
                "checksum": head_cs, "checksum_vs_oracle": head_verdict,
                "phases_ms": {"symbolic": s_ms, "numeric": k_ms},
                "roofline": {"bound": "hbm", "kernel": "numeric pass (tile multiply-accumulate), rank 0's launch", "achieved": ach,
                             "peak": peak, "unit": "GB/s", "frac": ach / peak,
                             "traffic": traffic, "traffic_source": traffic_src,
                             "algorithmic_bytes": alg, "peak_source": peak_src,
                             "useful_u64_mac_per_s": mac_rate, "int_mac_peak": mac_peak,
                             "int_mac_frac": (mac_rate / mac_peak) if mac_peak else None,
                             "int_mac_peak_source": "profiles/r01_u64mac_peak.jsonl (tools/u64mac_peak.cu: the kernel's inner loop on "
                                                    "registers and shared memory only, best of its grid sizes; not a driver-measured peak)"},
                "roofline_symbolic": {"bound": "hbm", "achieved": sb / (s_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                                      "frac": sb / (s_ms * 1e-3) / 1e9 / peak, "algorithmic_bytes": sb,
                                      "bytes_with_pair_lists": sb + pl,
                                      "frac_with_pair_lists": (sb + pl) / (s_ms * 1e-3) / 1e9 / peak,
                                      "frac_of_nominal_8TBs": sb / (s_ms * 1e-3) / 1e9 / 8000.0},
                "gpu_launches": int(launches), "clocks": clocks,
            }
            if e2e:
                line["e2e"] = e2e
            if chain:
                line["chain"] = chain
            if rmat:
                line["rmat_panels"] = rmat
            if floats:
                line["float32_tiles"] = floats
            if world == 1 and not args.no_cpu:
                try:
                    cb = cpu_reference_measure(n, k, d, steps=1, warmup=0)
                    line["cpu_baseline"] = {kk: cb[kk] for kk in ("value", "unit", "cores", "kind", "sample")}
                except Exception as ex:  # the baseline is reported, never load-bearing
                    line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": str(ex)}
            real_stdout.write(json.dumps(line) + "\n")
            real_stdout.flush()
        if comm:
            comm.close()
        ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def keep_count(a, bounds, rank):
    return int(((a[3] >= bounds[rank]) & (a[3] < bounds[rank + 1])).sum())


if __name__ == "__main__":
    sys.exit(main())
