"""numpy model of kernels_count.cuh (run-aggregated slot assignment in an arbitrary warp order + rank by source index):
checks that the result is the stable sort for sorted-ish and random key arrays. CPU only; not part of the product."""
import numpy as np


def clz32(x):
    return 32 - int(x).bit_length()


def model(keys, num_cells, rng):
    n = keys.shape[0]
    kc = np.minimum(keys, num_cells - 1)
    count = np.bincount(kc, minlength=num_cells)
    lb = np.concatenate([[0], np.cumsum(count)]).astype(np.int64)
    cursor = lb[:-1].copy()
    slot_keys = np.full(n, -1, np.int64)
    slot_idx = np.full(n, -1, np.int64)
    nwarps = (n + 31) // 32
    for w in rng.permutation(nwarps):  # atomics of different warps land in any order
        lanes = np.arange(32)
        i = w * 32 + lanes
        valid = i < n
        d = np.where(valid, kc[np.minimum(i, n - 1)], 0xffffffff)
        prev = np.concatenate([[d[0]], d[:-1]])
        heads = 0
        for l in range(32):
            if l == 0 or d[l] != prev[l]:
                heads |= 1 << l
        base = np.zeros(32, np.int64)
        head_of = np.zeros(32, np.int64)
        for l in rng.permutation(32):  # lanes of one instruction: any order too
            upto = ((2 << int(l)) - 1) & 0xffffffff
            head = 31 - clz32(heads & upto)
            head_of[l] = head
            if valid[l] and head == l:
                above = int(heads) & ~upto & 0xffffffff
                nxt = (above & -above).bit_length() - 1 if above else 32
                base[l] = cursor[d[l]]
                cursor[d[l]] += nxt - l
        for l in range(32):
            if valid[l]:
                slot = base[head_of[l]] + (l - head_of[l])
                assert slot_keys[slot] == -1
                slot_keys[slot] = keys[i[l]]
                slot_idx[slot] = i[l]
    assert (slot_keys >= 0).all()
    out_idx = np.full(n, -1, np.int64)
    for q in range(n):
        c = min(slot_keys[q], num_cells - 1)
        lo, hi = lb[c], lb[c + 1]
        rank = int((slot_idx[lo:hi] < slot_idx[q]).sum())
        assert out_idx[lo + rank] == -1
        out_idx[lo + rank] = slot_idx[q]
    return slot_keys, out_idx, lb


def main():
    rng = np.random.default_rng(7)
    for n, cells, kind in [(1000, 50, "sortedish"), (5000, 300, "random"), (777, 1, "one"), (4097, 4000, "sortedish"), (33, 7, "random")]:
        if kind == "random":
            keys = rng.integers(0, cells, n)
        elif kind == "one":
            keys = np.zeros(n, np.int64)
        else:
            keys = np.sort(rng.integers(0, cells, n))
            mv = rng.random(n) < 0.1
            keys = np.clip(keys + mv * rng.integers(-3, 4, n), 0, cells - 1)
        sk, oi, lb = model(keys.astype(np.int64), cells, rng)
        want = np.argsort(keys, kind="stable")
        assert np.array_equal(oi, want), (n, cells, kind)
        assert np.array_equal(sk, keys[want])
        print("ok", n, cells, kind)


if __name__ == "__main__":
    main()
