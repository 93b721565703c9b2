"""Writes a golden case's matrices as a reference-format folder, blocks in FILE order
(duplicates / unsorted keys preserved).  Shared by the CLI tests."""
import os


def write_raw_folder(folder, mats, k):
    os.makedirs(folder, exist_ok=True)
    with open(os.path.join(folder, "size"), "w") as f:
        f.write("%d %d\n" % (len(mats), k))
    for i, (rows, cols, br, bc, v) in enumerate(mats):
        v = v.reshape(-1, k, k)
        with open(os.path.join(folder, "matrix%d" % (i + 1)), "w") as f:
            f.write("%d %d\n%d\n" % (rows, cols, len(br)))
            for r, c, t in zip(br, bc, v):
                f.write("%d %d\n" % (r, c))
                for row in t:
                    f.write(" ".join(str(int(x)) for x in row) + "\n")
