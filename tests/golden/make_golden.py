"""Regenerates tests/golden/ from the mounted reference (run in the build container only).

1. copies the reference's own test fixtures (data files, not sources) that pin the hot path:
   data/input/{csv1,csv2}_{0..3}.csv, multi_join{1,2}.csv and the per-(world, rank) expected outputs
   data/output/{join_inner,multi_join,union,subtract,intersect}_{W}_{r}.csv
   (cpp/test/join_test.cpp:20-53, cpp/test/set_op_test.cpp:22-79), the unique fixture
   cpp/src/tutorial/data/duplicate_data_0.csv (cpp/test/table_op_test.cpp:49-86) and
   user_usage_tm_{1..4}.csv (python/test/test_dist_rl.py:76-100);
2. writes known_answers.json: MurmurHash3_x86_32 vectors produced by the reference's own
   util/murmur3.cpp (compiled by oracle/build_ref.sh into oracle/_ref/libref_murmur3.so), plus the
   expected row counts asserted by python/test/test_dist_rl.py.
"""
import ctypes
import json
import os
import shutil
import struct
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"


def main():
    dst = os.path.join(HERE, "ref_data")
    os.makedirs(dst, exist_ok=True)
    files = []
    for r in range(4):
        files += ["data/input/csv1_%d.csv" % r, "data/input/csv2_%d.csv" % r]
    files += ["data/input/multi_join1.csv", "data/input/multi_join2.csv"]
    for op in ("join_inner", "multi_join", "union", "subtract", "intersect"):
        for w in (1, 2, 4):
            for r in range(w):
                files.append("data/output/%s_%d_%d.csv" % (op, w, r))
    files.append("cpp/src/tutorial/data/duplicate_data_0.csv")
    files += ["cpp/src/tutorial/data/user_usage_tm_%d.csv" % i for i in range(1, 5)]
    for f in files:
        shutil.copyfile(os.path.join(REF, f), os.path.join(dst, os.path.basename(f)))

    subprocess.check_call([os.path.join(ROOT, "oracle", "build_ref.sh")])
    ref = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_murmur3.so"))
    ref.ref_murmur3_x86_32.restype = ctypes.c_uint32
    ref.ref_murmur3_x86_32.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_uint32]

    vectors = []
    samples = [b"", b"a", b"ab", b"abc", b"abcd", b"abcde", b"hello world", b"The quick brown fox jumps over the lazy dog",
               bytes(range(256))]
    for d in (10.0, 11.0, 21.0, 0.394, 1.0, 0.0, -1.5, 1e300, 2.2250738585072014e-308):
        samples.append(struct.pack("<d", d))
    for f in (1.5, 0.0, -2.25, 3.4e38):
        samples.append(struct.pack("<f", f))
    for i in range(64):
        samples.append(bytes((i * 37 + j * 11) & 0xFF for j in range(i)))
    for s in samples:
        for seed in (0, 1, 0x9747b28c):
            vectors.append({"hex": s.hex(), "seed": seed, "hash": int(ref.ref_murmur3_x86_32(s, len(s), seed))})

    known = {
        "generator": "tests/golden/make_golden.py over /root/reference (cylon 0.4.0)",
        "murmur3_x86_32": vectors,
        # python/test/test_dist_rl.py:76-100 (4 ranks, user_usage_tm_{rank+1}.csv joined with itself on col 0)
        "test_dist_rl": {"world": 4,
                         "join_rows": [1424, 1648, 2704, 1552], "join_cols": 8,
                         "union_rows": [62, 53, 53, 72], "intersect_rows": [62, 53, 53, 72],
                         "subtract_rows": [0, 0, 0, 0]},
        # cpp/test/table_op_test.cpp:73 -- column 3 of Unique(duplicate_data_0, cols={0}) after sorting
        "unique_col3": [1, 2, 3, 4, 5, 7, 10, 12, 13, 14, 15],
    }
    with open(os.path.join(HERE, "known_answers.json"), "w") as f:
        json.dump(known, f, indent=0)
    print("wrote", len(files), "fixtures and", len(vectors), "murmur vectors")


if __name__ == "__main__":
    main()
