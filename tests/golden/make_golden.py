"""Generates tests/golden/ref_archives.npz: archives produced by the REFERENCE's own CUDA build
(oracle/_ref/libdietgpu_ref.so, compiled from /root/reference by oracle/build_ref.sh) on a B200.

Run on the GPU box (the reference has no CPU codec, so its output can only be captured there):

    gpurun -- 'python tests/golden/make_golden.py gpurun_out/ref_archives.npz'

then copy gpurun_out/ref_archives.npz to tests/golden/. Inputs are not stored: every case is
regenerated from its seed by tests/datagen.py, so the fixture holds only the case table and the
reference's archive bytes (output buffers pre-filled with 0xCD so that bytes the reference never
writes are visible as such). tests/test_golden.py checks the CPU oracle against these bytes
under the defined-byte mask on every CPU run; the GPU tests check our kernels against them too.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

# (kind, float type or 0, n, seed, prob_bits, checksum, extra)   extra: lambda for ans, sparsity for sparse
CASES = [
    ("ans", 0, 1, 11, 10, 0, 20.0),
    ("ans", 0, 31, 12, 10, 0, 20.0),
    ("ans", 0, 4096, 13, 10, 0, 20.0),
    ("ans", 0, 4097, 14, 10, 1, 20.0),
    ("ans", 0, 70001, 15, 10, 0, 20.0),
    ("ans", 0, 70001, 16, 9, 0, 5.0),
    ("ans", 0, 70001, 17, 11, 1, 100.0),
    ("ans", 0, 262144, 18, 10, 0, 20.0),
    ("float", 1, 1, 21, 10, 0, 1.0),
    ("float", 1, 5000, 22, 10, 0, 1.0),
    ("float", 1, 65537, 23, 10, 1, 1.0),
    ("float", 2, 33, 24, 10, 0, 0.02),
    ("float", 2, 4096, 25, 10, 0, 0.02),
    ("float", 2, 100000, 26, 10, 0, 0.02),
    ("float", 2, 100001, 27, 9, 1, 1.0),
    ("float", 3, 7, 28, 10, 0, 1.0),
    ("float", 3, 50000, 29, 10, 0, 1.0),
    ("float", 3, 131073, 30, 11, 1, 1.0),
    ("float", 4, 5, 31, 10, 0, 1.0),
    ("float", 4, 40000, 32, 10, 0, 1.0),
    ("float", 4, 65536, 33, 10, 1, 1.0),
    ("sparse", 3, 4096, 41, 10, 0, 0.9),
    ("sparse", 3, 100000, 42, 10, 0, 0.9),
    ("sparse", 3, 100003, 43, 10, 0, 0.5),
    ("sparse", 2, 50000, 44, 10, 0, 0.9),
    ("sparse", 1, 50001, 45, 10, 0, 0.99),
]


def case_input(case):
    """The seeded input of a case (u8 bytes for ans, unsigned words for float / sparse)."""
    import datagen
    kind, ft, n, seed, _pb, _ck, extra = case
    if kind == "ans":
        return datagen.exp_bytes(n, lam=extra, seed=seed)
    if kind == "float":
        return datagen.float_words(ft, n, seed=seed, scale=extra)
    x = datagen.float_words(ft, n, seed=seed, scale=1.0, sparsity=extra)
    # The reference's sparse compaction assumes word N-2 is non-zero; when it is zero its dense array holds
    # one uninitialised word and the nested float archive is not reproducible. All
    # sparse cases but one (seed 43) therefore pin the last two words to non-zero values.
    one = {1: 0x3C00, 2: 0x3F80, 3: 0x3F800000, 4: 0x3FF0000000000000}[ft]
    x[-1] = one + 1
    x[-2] = 0 if seed == 43 else one
    return x


def main(out_path):
    import gpu_harness as H
    assert H.ref_lib() is not None, "oracle/_ref/libdietgpu_ref.so is not built"
    blobs, sizes = [], []
    for case in CASES:
        kind, ft, n, _seed, pb, ck, _extra = case
        dev = [H.to_dev(case_input(case))]
        if kind == "ans":
            m, sz = H.ref_ans_encode(dev, pb, bool(ck))
        else:
            m, sz = H.ref_float_compress(ft, dev, [n], pb, bool(ck), sparse=(kind == "sparse"))
        blobs.append(H.row_bytes(m, 0, int(sz[0])))
        sizes.append(int(sz[0]))
    table = np.array([(k, ft, n, seed, pb, ck, ex) for k, ft, n, seed, pb, ck, ex in CASES],
                     dtype=[("kind", "U8"), ("ft", "i4"), ("n", "i8"), ("seed", "i4"), ("pb", "i4"),
                            ("ck", "i4"), ("extra", "f8")])
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    np.savez_compressed(out_path, cases=table, offsets=offs, data=np.concatenate(blobs))
    print(f"wrote {out_path}: {len(CASES)} cases, {offs[-1]} archive bytes")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "ref_archives.npz"))
