"""Host-side partitioning logic (CPU only)."""
import numpy as np
import pytest

import bsp_dist


@pytest.mark.parametrize("n,P,want", [
    (64, 8, [(8 * r, 8 * r + 7) for r in range(8)]),
    (15, 8, [(r, r) for r in range(7)] + [(7, 14)]),
    (5, 2, [(0, 1), (2, 4)]),
    (3, 8, [(0, 2)]),
    (7, 1, [(0, 6)]),
])
def test_split_matches_reference_rule(n, P, want):
    assert bsp_dist.split_chain(n, P) == want


def test_tree_levels():
    assert bsp_dist.tree_levels(5) == [[(0, 1), (2, 3)], [(0, 1)], [(0, 1)]]
    assert bsp_dist.tree_levels(1) == []


@pytest.mark.parametrize("P", range(1, 17))
@pytest.mark.parametrize("n", [1, 2, 5, 16, 23, 64])
def test_step_doubling_has_the_reference_association(n, P):
    assert bsp_dist.association(n, P) == bsp_dist.association_by_schedule(n, P)


def test_association_examples():
    assert bsp_dist.association(5, 1) == "(((M0M1)(M2M3))M4)"
    assert bsp_dist.association(5, 2) == "((M0M1)((M2M3)M4))"
    assert bsp_dist.association(7, 7) == "(((M0M1)(M2M3))((M4M5)M6))"


def test_balanced_panels():
    w = np.array([1, 1, 1, 1, 100, 1, 1, 1])
    b = bsp_dist.balanced_panels(w, 2)
    assert b[0] == 0 and b[-1] == 8 and b == sorted(b)
    loads = [w[b[i]:b[i + 1]].sum() for i in range(2)]
    assert max(loads) <= 104
    assert bsp_dist.balanced_panels(np.ones(10), 5) == [0, 2, 4, 6, 8, 10]


@pytest.mark.parametrize("n", list(range(1, 18)))
def test_panel_orders_cover_every_rank_once(n):
    """tree_merge_distributed: after every level the panels of a group's product are held by
    exactly the ranks of the group, each once; a rank hands rows only to rank+s inside its group."""
    levels, final = bsp_dist.panel_orders(n)
    assert sorted(final) == list(range(n)) and final[0] == 0
    assert [(s, g) for s, g, _, _ in levels] == bsp_dist.merge_groups(n)
    seen = {r: [r] for r in range(n)}
    for s, group, left, right in levels:
        g = group[0]
        assert left == seen[g] and right == seen[g + s]
        assert sorted(left + right) == group
        assert all(q - s in left for q in right)       # every right-hand rank has a partner that feeds it
        new = []
        for p in left:
            new.append(p)
            if p + s in right:
                new.append(p + s)
        seen[g] = new
        del seen[g + s]
    assert seen == {0: final}
