#!/usr/bin/env python3
"""Pin the commit of a trace that does not fit host RAM (the 2^23 x 665 headline) against the COMPILED REFERENCE.

Runs ONLY in the build container (needs oracle/_ref/libzkref.so = the unmodified reference sources, oracle/Makefile).
SURVEY.md section 8(d) "parity check", option (iii): the LDE is independent per column (ntt_goldilocks.cpp:103-110), so the
extended trace is produced strip by strip with the reference's own NTT_Goldilocks::extendPol; the row sponge is advanced
strip by strip with the reference's own hash_full_result (zkref_sponge_absorb = the loop body of linear_hash,
poseidon_goldilocks.cpp:240-256, "capacity <- previous out[0..3]"); the node levels are the reference's merkletree over
the 4-word digests (rows of <= 4 words are copied, not hashed, :234-239, so its leaf level IS the digests).

Writes tests/golden/headline_pin.json: the root, sampled leaf digests, sampled tree nodes of every level and sampled full
rows of the extended trace.  tests/test_gpu_parity.py::test_headline_against_reference_pin and bench.py's KNOWN_ROOTS read it.

  python tools/pin_headline_root.py [--log-rows 23 --cols 665 --log-blowup 1 --strip 40] [--check]
--check additionally runs the whole-matrix reference (extendPol + merkletree) and compares every word (small shapes only).
"""
import argparse, ctypes, json, pathlib, sys, time
import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
SEED = 0x9E3779B97F4A7C15
u64 = ctypes.c_uint64
vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log-rows", type=int, default=23)
    ap.add_argument("--cols", type=int, default=665)
    ap.add_argument("--log-blowup", type=int, default=1)
    ap.add_argument("--strip", type=int, default=40, help="columns per strip (multiple of 8)")
    ap.add_argument("--rows-sampled", type=int, default=24)
    ap.add_argument("--leaves-sampled", type=int, default=1024)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--out", default=str(ROOT / "tests" / "golden" / "headline_pin.json"))
    a = ap.parse_args()
    assert a.strip % 8 == 0
    R = ctypes.CDLL(str(ROOT / "oracle" / "_ref" / "libzkref.so"))
    O = ctypes.CDLL(str(ROOT / "oracle" / "liboracle.so"))
    R.zkref_threads.restype = ctypes.c_int
    n, b, c = 1 << a.log_rows, 1 << a.log_blowup, a.cols
    n_ext = n * b
    rng = np.random.default_rng(20261105)
    rows_s = np.unique(np.concatenate([[0, 1, n_ext // 2, n_ext - 1], rng.integers(0, n_ext, a.rows_sampled)]))[: a.rows_sampled + 4]
    leaves_s = np.unique(np.concatenate([[0, n_ext - 1], rng.integers(0, n_ext, a.leaves_sampled)]))
    sampled_rows = np.zeros((rows_s.size, c), dtype=np.uint64)
    cap = np.zeros(n_ext * 4, dtype=np.uint64)
    t0 = time.time()
    for c0 in range(0, c, a.strip):
        w = min(a.strip, c - c0)
        inp = np.empty(n * w, dtype=np.uint64)
        O.oracle_fill_block(vp(inp), u64(n), u64(c0), u64(w), u64(c), u64(SEED))
        out = np.zeros(n_ext * w, dtype=np.uint64)
        buf = np.zeros(n_ext * w, dtype=np.uint64)
        R.zkref_extend_pol(vp(out), vp(inp), u64(n_ext), u64(n), u64(w), vp(buf))
        del buf, inp
        R.zkref_sponge_absorb(vp(cap), vp(out), u64(w), u64(w), u64(n_ext), int(c0 == 0))
        R.zkref_canon(vp(out), u64(out.size))
        sampled_rows[:, c0:c0 + w] = out.reshape(n_ext, w)[rows_s]
        del out
        print("strip %d..%d done, %.0f s" % (c0, c0 + w, time.time() - t0), file=sys.stderr, flush=True)
    tree = np.zeros((2 * n_ext - 1) * 4, dtype=np.uint64)
    R.zkref_merkletree(vp(tree), vp(cap), u64(4), u64(n_ext), u64(1))
    R.zkref_canon(vp(tree), u64(tree.size))
    hexw = lambda v: ["%016x" % int(x) for x in np.asarray(v).reshape(-1)]
    levels, off, cnt, lvl = [], 0, n_ext, 0
    while True:
        idx = np.unique(np.concatenate([[0, cnt - 1], rng.integers(0, cnt, 6)])) if lvl else leaves_s
        levels.append({"level": lvl, "nodes": cnt, "index": [int(i) for i in idx],
                       "digests": [hexw(tree[off + 4 * int(i): off + 4 * int(i) + 4]) for i in idx]})
        if cnt == 1:
            break
        off += cnt * 4
        cnt //= 2
        lvl += 1
    doc = {"generator": "tools/pin_headline_root.py (compiled unmodified reference, strip-wise; SURVEY 8(d) option iii)",
           "input": "in[i*C + j] = splitmix64(seed + i*C + j) mod p, seed 0x9E3779B97F4A7C15",
           "log_rows": a.log_rows, "cols": c, "log_blowup": a.log_blowup, "strip": a.strip, "reference_threads": R.zkref_threads(),
           "root": hexw(tree[-4:]), "levels": levels,
           "rows": {"index": [int(i) for i in rows_s], "words": [hexw(r) for r in sampled_rows]},
           "wall_s": time.time() - t0}
    if a.check:
        inp = np.empty(n * c, dtype=np.uint64)
        O.oracle_fill(vp(inp), u64(n * c), u64(SEED), u64(0))
        out = np.zeros(n_ext * c, dtype=np.uint64)
        buf = np.zeros(n_ext * c, dtype=np.uint64)
        R.zkref_extend_pol(vp(out), vp(inp), u64(n_ext), u64(n), u64(c), vp(buf))
        t2 = np.zeros_like(tree)
        R.zkref_merkletree(vp(t2), vp(out), u64(c), u64(n_ext), u64(1))
        R.zkref_canon(vp(t2), u64(t2.size))
        R.zkref_canon(vp(out), u64(out.size))
        assert np.array_equal(t2, tree), "strip-wise tree differs from the whole-matrix reference"
        assert np.array_equal(out.reshape(n_ext, c)[rows_s], sampled_rows)
        print("check ok: strip-wise == whole-matrix reference (%d tree words)" % tree.size, file=sys.stderr)
    pathlib.Path(a.out).write_text(json.dumps(doc, indent=0) + "\n")
    print("root", " ".join(doc["root"]), "-> %s (%.0f s)" % (a.out, doc["wall_s"]))


if __name__ == "__main__":
    main()
