"""CPU model of the fp16 filter's class-minima epilogue (orange3-spark_b200/csrc/kmeans_tcb.cu, TB_GRID).

The kernel never keeps a running (best, second best) over the k scores of a row.  Every score falls into two classes
-- its M-class (the 16 consecutive centres around it) and its A-class (its column mod 16) -- and only the three smallest
class minima of either kind reach the finish warps.  What the kernel relies on, restated here in numpy on random and
adversarial integer scores and checked against a full sort:

  * best = smallest M-minimum = smallest A-minimum;
  * second best = min(second smallest M-minimum, second smallest A-minimum)  (two centres never share both classes);
  * every centre whose score is below T lies in (M-classes with minimum < T) x (A-classes with minimum < T), so when at
    most two classes of either kind are below T, the <= 4 enumerated candidates contain every such centre;
  * with the work split between two epilogue warps (alternate centre tiles of 128), the same holds per warp and the
    candidate sets simply add up.

No GPU, no library call: this is the arithmetic argument of DESIGN.md 4.1, executable."""
import numpy as np
import pytest


def class_minima(scores):
    """scores: int array [k], k % 16 == 0.  Returns (sorted M-minima with their class ids, sorted A-minima with ids)."""
    k = len(scores)
    m = scores.reshape(k // 16, 16)
    mmin = m.min(axis=1)                       # one per group of 16 consecutive centres
    amin = m.min(axis=0)                       # one per column mod 16
    mo, ao = np.argsort(mmin, kind="stable"), np.argsort(amin, kind="stable")
    return (mmin[mo], mo), (amin[ao], ao)


def enumerate_row(per_warp, thr):
    """per_warp: list of (scores, global centre ids) per epilogue warp.  Mirrors the finish warps' logic."""
    best = min(int(s.min()) for s, _ in per_warp)
    T = best + thr
    total, many, cand = 0, False, []
    for s, ids in per_warp:
        (mv, mi), (av, ai) = class_minima(s)
        nM = int((mv[:3] <= T).sum())
        nA = int((av[:3] <= T).sum())
        if nM == 3 or nA == 3:
            many = True
        nM, nA = min(nM, 2), min(nA, 2)
        total += nM * nA
        for a in range(nM):
            for b in range(nA):
                cand.append(int(ids[mi[a] * 16 + ai[b]]))
    return best, total, many, cand


def split_two_warps(scores):
    """Alternate centre tiles of 128 go to the two epilogue warps of a lane quarter."""
    k = len(scores)
    tiles = np.arange(k).reshape(k // 128, 128)
    out = []
    for h in (0, 1):
        ids = tiles[h::2].ravel()
        if len(ids):
            out.append((scores[ids], ids))
    return out


@pytest.mark.parametrize("k", [128, 256, 384, 1024, 2048])
def test_second_best_is_the_smaller_of_the_two_second_class_minima(k):
    rng = np.random.default_rng(k)
    for trial in range(300):
        spread = [4, 50, 10_000][trial % 3]  # many ties / some ties / none
        s = rng.integers(0, spread * k, size=k)
        (mv, _), (av, _) = class_minima(s)
        srt = np.sort(s)
        assert mv[0] == av[0] == srt[0]
        assert min(mv[1], av[1]) == srt[1]


@pytest.mark.parametrize("k", [128, 256, 640, 1024])
def test_candidate_enumeration_contains_every_centre_within_the_threshold(k):
    rng = np.random.default_rng(1000 + k)
    n_cert = n_cand = n_many = 0
    for trial in range(600):
        s = rng.integers(0, 200_000, size=k)
        # plant one to four near-competitors of the best at random places (same group, same column, or neither)
        b = int(rng.integers(k))
        s[b] = -1000
        thr = int(rng.integers(1, 40))
        for _ in range(int(rng.integers(0, 4))):
            kind = rng.integers(3)
            j = {0: (b // 16) * 16 + int(rng.integers(16)),           # same M-class
                 1: (int(rng.integers(k // 16)) * 16 + b % 16),      # same A-class
                 2: int(rng.integers(k))}[int(kind)]
            s[j] = -1000 + int(rng.integers(0, 2 * thr))
        for per_warp in ([(s, np.arange(k))], split_two_warps(s)):   # one warp sees all (TB_SKEW) / two warps share
            best, total, many, cand = enumerate_row(per_warp, thr)
            within = set(np.flatnonzero(s <= best + thr).tolist())
            assert best == s.min()
            if many:
                n_many += 1
                continue  # the kernel sends such rows to the 3xTF32 re-check
            assert total == len(cand) and len(set(cand)) == len(cand)
            assert within <= set(cand), (sorted(within), sorted(cand))
            if total == 1:
                n_cert += 1
                assert within == {int(np.argmin(s))}  # certain: nobody else within the threshold
            elif total <= 4:
                n_cand += 1
    assert n_cert > 50 and n_cand > 50 and n_many > 10  # every branch of the finish logic was exercised


def test_a_hidden_competitor_is_visible_in_the_other_class():
    """The case the earlier pair check could not see: best and second best in ONE group of 16 (the M-minimum hides the
    second) -- the A-minima show it; and both in ONE column class across groups -- the M-minima show it."""
    k = 256
    s = np.full(k, 1_000_000)
    s[37], s[41] = 10, 12          # same group (32..47), different columns
    (mv, _), (av, _) = class_minima(s)
    assert mv[1] == 1_000_000 and av[1] == 12
    s = np.full(k, 1_000_000)
    s[37], s[37 + 64] = 10, 12     # same column (5), different groups
    (mv, _), (av, _) = class_minima(s)
    assert mv[1] == 12 and av[1] == 1_000_000
