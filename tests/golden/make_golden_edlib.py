#!/usr/bin/env python3
"""Golden vectors of edlib's NW alignment path, made by the REFERENCE's own edlib (oracle/_ref/libedlib_ref.so =
/root/reference/edlib/edlib.cpp compiled as it lies; `make -C oracle ref`): edlibAlign(EDLIB_MODE_NW, EDLIB_TASK_PATH) on
seeded pairs -> tests/golden/edlib_paths.npz (sequences, k, distance, alignment ops).  Run where /root/reference exists:
    python tests/golden/make_golden_edlib.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib                                   # noqa: E402
from test_oracle_vs_ref import _pairs_for_alignment  # noqa: E402


def main():
    rng = np.random.default_rng(2024)
    qs, ts, ks, ds, ops_all = [], [], [], [], []
    for i, (q, t) in enumerate(_pairs_for_alignment(rng, 120)):
        k = -1 if i % 3 else 10
        d, ops = oracle_lib.edlib_nw_ops(q.tobytes(), t.tobytes(), k)
        qs.append(q)
        ts.append(t)
        ks.append(k)
        ds.append(d)
        ops_all.append(ops if ops is not None else np.zeros(0, np.uint8))
    np.savez_compressed(os.path.join(HERE, "edlib_paths.npz"),
                        q=np.concatenate(qs), qlen=np.array([x.size for x in qs], np.int32),
                        t=np.concatenate(ts), tlen=np.array([x.size for x in ts], np.int32),
                        k=np.array(ks, np.int32), dist=np.array(ds, np.int32),
                        ops=np.concatenate(ops_all), n_ops=np.array([x.size for x in ops_all], np.int32))
    print("edlib_paths.npz:", len(qs), "pairs,", int(sum(x.size for x in ops_all)), "ops,", sum(1 for d in ds if d < 0), "beyond k")


if __name__ == "__main__":
    main()
