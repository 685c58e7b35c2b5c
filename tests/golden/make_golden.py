"""Copies the reference's own golden fixtures for the hot path into tests/golden/ (run in the dev
container, where /root/reference exists). Source: hnswlib/tests/cpp/data/{batch_final,queries,gt}.bin +
config.txt, generated upstream by hnswlib/tests/cpp/update_gen_data.py (np.random.seed(1), 100000 x 8
L2-normalised uniform vectors, 1000 queries, gt = exact top-5 ids by dot product); used by
hnswlib/tests/cpp/updates_test.cpp:149-172. Also writes spaces_known_answers.json restating
hnswlib/tests/python/bindings_test_spaces.py:11-33."""
import json
import os
import shutil

REF = "/root/reference/hnswlib/tests/cpp/data"
HERE = os.path.dirname(os.path.abspath(__file__))

if __name__ == "__main__":
    for f in ("batch_final.bin", "queries.bin", "gt.bin", "config.txt"):
        shutil.copyfile(os.path.join(REF, f), os.path.join(HERE, "hnswlib_updates_" + f))
    json.dump({
        "source": "hnswlib/tests/python/bindings_test_spaces.py:11-33",
        "data1": [[1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 0, 1], [1, 1, 1]],
        "query_row": 4,
        "l2": [0., 1., 2., 2., 2.],
        "ip": [-2., -1., 0., 0., 0.],
        "rightdims": list(range(1, 128, 3)),
        "leftdims": list(range(1, 32, 5)),
        "tol_mean_abs": 1e-3,
    }, open(os.path.join(HERE, "spaces_known_answers.json"), "w"), indent=1)
