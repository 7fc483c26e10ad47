"""Generates tests/golden/eval_golden.npz from the reference's OWN evaluator
(Registration/registration_dataset/evaluate_rt.py, imported with an empty stand-in for the absent open3d, which
the scoring functions never touch) on the two reg_result.txt files the reference ships.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_eval.py
"""
import contextlib
import io
import os
import sys
import types

import numpy as np

REF = "/root/reference/Registration"
HERE = os.path.dirname(os.path.abspath(__file__))


def rows(path):
    return np.array([[float(x) for x in line.split(",")] for line in open(path).readlines()[1:]])


def main():
    sys.modules.setdefault("open3d", types.ModuleType("open3d"))
    sys.path.insert(0, os.path.join(REF, "registration_dataset"))
    import evaluate_rt as ref
    a = os.path.join(REF, "reg_result.txt")                            # "# idx1,..." header (np.savetxt)
    b = os.path.join(REF, "registration_dataset", "reg_result.txt")    # plain header
    out = {"rows_a": rows(a), "rows_b": rows(b), "header_a": open(a).readline().rstrip("\n"),
           "header_b": open(b).readline().rstrip("\n")}
    with contextlib.redirect_stdout(io.StringIO()):
        out["eval_ab"] = np.array(ref.evaluate_rt(a, b))
        out["eval_ba"] = np.array(ref.evaluate_rt(b, a))
        out["eval_aa"] = np.array(ref.evaluate_rt(a, a))
    ra, rb = ref.read_reg_results(a), ref.read_reg_results(b)
    diffs = []
    for i in range(1, 40):
        _, _, ta, qa = ref.reg_result_row_to_array(ra[i])
        _, _, tb, qb = ref.reg_result_row_to_array(rb[i])
        diffs.append(ref.get_P_diff(ref.get_P_from_Rt(qa.as_matrix(), ta), ref.get_P_from_Rt(qb.as_matrix(), tb)))
    out["p_diff_first39"] = np.array(diffs)
    np.savez_compressed(os.path.join(HERE, "eval_golden.npz"), **out)
    print({k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items()}, out["eval_ab"], out["eval_aa"])


if __name__ == "__main__":
    main()
