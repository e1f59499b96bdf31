"""The reach truncation of the beam fringe (csrc/beam.cu, reach_keep) as a model in plain Python: a fringe that stores
only what the remaining pop rounds can still reach must pop EXACTLY what the reference's list pops
(seq2seq.py:1397-1420, :1515-1532), round by round, and account for the same list length.

CPU only; the CUDA kernels are checked against the oracle's statistics in tests/test_gpu_beam.py.  This model follows
the kernel's rule line by line (the three bounds of reach_keep) so that the argument in its comment is executable.
"""
import random

import pytest


def _children(rng_seed, node, n_max, p_final, spread):
    """Deterministic function of the parent: [(key, id, final)] -- both fringes must see the same children."""
    r = random.Random(rng_seed * 1000003 + node[1])
    out = []
    for q in range(r.randint(0, n_max)):
        out.append((node[0] + r.random() * spread, node[1] * 17 + q + 1, r.random() < p_final))
    return out


def _pop(fringe, batch):
    """Prefix with at most `batch` open nodes (finals inside it are consumed, :1400-1415)."""
    taken, opened = [], 0
    i = 0
    while i < len(fringe) and opened < batch:
        e = fringe[i]
        taken.append(e)
        if not e[2]:
            opened += 1
        i += 1
    return taken, fringe[i:]


def _run(seed, T, batch, n_max, p_final, spread, truncate):
    rounds_max = 2 * T
    keep_ref = rounds_max * batch
    fringe = [(0.0, 1, False)]
    gone, fin_in = 0, 0
    log = []
    it = 0
    while it < rounds_max:
        taken, rest = _pop(fringe, batch)
        opened = [e for e in taken if not e[2]]
        if not opened:
            break
        fin_in -= sum(1 for e in taken if e[2])
        it += 1
        new = sorted(c for e in opened for c in _children(seed, e, n_max, p_final, spread))
        merged = sorted(rest + new)
        listed = min(len(rest) + gone + len(new), keep_ref)
        q = min(keep_ref, len(merged))
        if truncate:
            if gone > 0:
                q = min(q, (len(rest) + sum(1 for c in new if c < rest[-1])) if rest else 0)
            new_finals = sum(1 for c in new if c[2])
            q = min(q, (rounds_max - it) * batch + fin_in + new_finals + 1)
            fin_in += new_finals
        nxt = merged[:q] if q < len(merged) else merged
        # what the kernel may look at behind the pop: the next entry (near-tie statistic)
        log.append((tuple(e[1] for e in taken), rest[0][1] if rest else None, listed))
        fringe, gone = nxt, listed - len(nxt)
    log.append(("left", len(fringe) + gone))
    return log, max(len(fringe), 1)


@pytest.mark.parametrize("seed", range(12))
@pytest.mark.parametrize("T,batch,n_max,p_final,spread", [(6, 4, 5, 0.1, 1.0), (10, 8, 6, 0.3, 0.2), (4, 3, 9, 0.0, 5.0),
                                                            (12, 4, 3, 0.6, 1.0)])
def test_truncated_fringe_pops_what_the_full_list_pops(seed, T, batch, n_max, p_final, spread):
    full, _ = _run(seed, T, batch, n_max, p_final, spread, truncate=False)
    cut, _ = _run(seed, T, batch, n_max, p_final, spread, truncate=True)
    assert full == cut


def test_truncation_actually_shortens_a_saturated_fringe():
    """With many children per node the list sits at its cap; the stored prefix must shrink with the rounds left."""
    T, batch = 10, 8
    _, last_full = _run(3, T, batch, 9, 0.05, 1.0, truncate=False)
    _, last_cut = _run(3, T, batch, 9, 0.05, 1.0, truncate=True)
    assert last_full > 4 * last_cut, (last_full, last_cut)
