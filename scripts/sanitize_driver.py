"""Small workload for compute-sanitizer (scripts/sanitize.sh): every kernel of the library once or twice, both compress
schedules, sizes small enough for the tools' 10-100 x slowdown. Results are checked against the CPU oracle."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import datagen  # noqa: E402
import gpu_harness as H  # noqa: E402
import oracle  # noqa: E402
from dietgpu_fork_b200 import _capi  # noqa: E402


def main():
    for sched, tiles in (("fused", 0), ("separate", 1 << 40)):
        _capi.lib.dgb_set_fused_min_tiles(tiles)
        for ft in (1, 2, 3, 4):
            counts = [70001, 4096, 33, 0, 140000]
            host = [datagen.float_words(ft, n, seed=7 + i) for i, n in enumerate(counts)]
            dev = [H.to_dev(h) for h in host]
            m, sz = H.ours_float_compress(ft, dev, counts, 10, checksum=(ft == 2))
            for i, h in enumerate(host):
                exp = oracle.float_compress(ft, h, 10, use_checksum=(ft == 2))
                assert np.array_equal(H.row_bytes(m, i, sz[i]), exp), (sched, ft, i)
            outs, ok, _ = H.ours_float_decompress(ft, [m[i] for i in range(len(counts))], counts, 10, checksum=(ft == 2))
            assert ok.all() and all(np.array_equal(o.cpu().numpy(), h.view(np.uint8)) for o, h in zip(outs, host))
        sizes = [1, 4097, 100000]
        hb = [datagen.exp_bytes(n, 20.0, seed=3 + i) for i, n in enumerate(sizes)]
        db = [H.to_dev(h) for h in hb]
        m, sz = H.ours_ans_encode(db, 10, checksum=True)
        for i, h in enumerate(hb):
            assert np.array_equal(H.row_bytes(m, i, sz[i]), oracle.ans_encode(h, 10, use_checksum=True))
        outs, ok, _ = H.ours_ans_decode([m[i] for i in range(3)], sizes, 10, checksum=True)
        assert ok.all() and all(np.array_equal(o.cpu().numpy(), h) for o, h in zip(outs, hb))
        hist = torch.from_numpy(np.stack([np.bincount(h, minlength=256) for h in hb]).astype(np.int32)).cuda()
        H.ours_ans_encode(db, 10, histogram=hist)  # k_normalize
    _capi.lib.dgb_set_fused_min_tiles(-1)
    for ft in (2, 3):  # sparse, both directions
        counts = [8191, 100000]
        host = [datagen.float_words(ft, n, seed=30 + i, sparsity=0.9) for i, n in enumerate(counts)]
        dev = [H.to_dev(h) for h in host]
        m, sz = H.ours_float_compress(ft, dev, counts, 10, sparse=True)
        for i, h in enumerate(host):
            assert np.array_equal(H.row_bytes(m, i, sz[i]), oracle.sparse_float_compress(ft, h, 10))
        outs, ok, _ = H.ours_float_decompress(ft, [m[i] for i in range(2)], counts, 10, sparse=True)
        assert all(np.array_equal(o.cpu().numpy(), h.view(np.uint8)) for o, h in zip(outs, host))
    torch.cuda.synchronize()
    print("sanitize_driver ok")


if __name__ == "__main__":
    main()
