"""Generates tests/golden/check_statistics.json with the REFERENCE's own TemplatedCheckStatistics<T>
(duckdb/src/planner/filter/bloom_filter.cpp:138-196, compiled unmodified into oracle/_ref/librpt_ref.so by
oracle/Makefile).  Run in the container where /root/reference is mounted:

    python tests/golden/make_check_statistics.py

The cases are the ones tests/test_oracle_cpu.py::test_check_statistics_pinned_to_reference replays (same seeds);
the frozen outputs let the pin hold on machines where the reference cannot be compiled (the GPU box)."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

import oracle_lib  # noqa: E402
import test_oracle_cpu as T  # noqa: E402


def main():
    ref = oracle_lib.ref()
    assert ref is not None, "needs oracle/_ref/librpt_ref.so (make -C oracle ref)"
    oracle = oracle_lib.oracle()
    out = {}
    for type_id in sorted(T.INT_TYPES):
        rng = np.random.default_rng(1000 + type_id)
        entry = {}
        for sparse in (False, True):
            words = T.stats_filter(oracle, type_id, sparse)
            # the filter itself is re-made by the reference's InsertHashes and must agree
            pairs = T.stats_cases(type_id, rng, n_random=60)
            mins, maxs = T.stats_arrays(type_id, pairs)
            has = np.ones(len(pairs), dtype=np.uint8)
            has[3] = 0
            entry["sparse" if sparse else "dense"] = ref.check_statistics(words, type_id, mins, maxs, has).tolist()
        out[str(type_id)] = entry
    path = os.path.join(HERE, "check_statistics.json")
    with open(path, "w") as f:
        json.dump(out, f)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
