"""Host utility mirroring Registration/registration_dataset/evaluate_rt.py: scores a
reg_result.txt against ground truth, so that C2 runs written by `icp.register_pairs` can be checked without the
reference tree.  Same names, arguments and return values; file handling and 342 x 9 arithmetic only (nothing
here is on the GPU path, and nothing here needs Open3D).

Reference behaviour kept on purpose:
  * rows are split on ',' and the FIRST line (the header, with or without numpy's "# " prefix) is skipped
    (evaluate_rt.py:92); the success rate nevertheless divides by the number of LINES including it (:114);
  * a pair succeeds when RTE < 2.0 and RRE < 5.0 (:16-18); RTE / RRE are averaged over the successes only;
  * P_diff = inv(P_pred) . P_gt; RTE = |t(P_diff)|; RRE = sum |euler xyz (degrees)| of R(P_diff)   (:21-29);
  * quaternions in the file are w, x, y, z (:65-70).
"""
import numpy as np
from scipy.spatial.transform import Rotation


def get_P_from_Rt(R, t):
    """evaluate_rt.py:9-13."""
    P = np.eye(4)
    P[:3, :3] = np.asarray(R, np.float64)
    P[:3, 3] = np.asarray(t, np.float64).reshape(3)
    return P


def get_P_diff(P_pred_np, P_gt_np):
    """evaluate_rt.py:21-29 -> (t_diff, angles_diff)."""
    D = np.linalg.inv(P_pred_np) @ P_gt_np
    rte = float(np.linalg.norm(D[:3, 3]))
    rre = float(np.abs(Rotation.from_matrix(D[:3, :3]).as_euler("xyz", degrees=True)).sum())
    return rte, rre


def is_registration_successful(P_pred_np, P_gt_np):
    """evaluate_rt.py:16-18 -> (ok, rte, rre)."""
    rte, rre = get_P_diff(P_pred_np, P_gt_np)
    return rte < 2.0 and rre < 5.0, rte, rre


def read_oxford_bin(bin_path):
    """evaluate_rt.py:45-51: 6 x N float32 [x, y, z, nx, ny, nz]."""
    flat = np.fromfile(bin_path, dtype=np.float32)
    return flat.reshape(-1, 6).T


def read_reg_results(file_path, splitter=","):
    """evaluate_rt.py:54-63: every line (header included) as a list of stripped strings."""
    with open(file_path, "r") as f:
        return [[item.strip() for item in line.split(splitter)] for line in f if line != ""]


def reg_result_row_to_array(reg_result_row):
    """evaluate_rt.py:65-75 -> (idx1, idx2, t (3,), scipy Rotation)."""
    idx1, idx2 = int(reg_result_row[0]), int(reg_result_row[1])
    vals = np.array([float(x) for x in reg_result_row[2:9]])
    w, x, y, z = vals[3:7]
    return idx1, idx2, vals[:3], Rotation.from_quat([x, y, z, w])


def evaluate_rt(gt_file_path, predict_file_path, verbose=False):
    """evaluate_rt.py:78-121 -> (success rate, mean RTE, mean RRE over the successes)."""
    gt, pred = read_reg_results(gt_file_path), read_reg_results(predict_file_path)
    if len(gt) != len(pred):
        raise AssertionError(f"{len(gt)} ground-truth lines vs {len(pred)} predicted lines")
    ok_n, rte_sum, rre_sum = 0, 0.0, 0.0
    for g_row, p_row in zip(gt[1:], pred[1:]):
        g1, g2, g_t, g_rot = reg_result_row_to_array(g_row)
        p1, p2, p_t, p_rot = reg_result_row_to_array(p_row)
        if (g1, g2) != (p1, p2):
            raise AssertionError(f"pair mismatch: ground truth {(g1, g2)} vs prediction {(p1, p2)}")
        ok, rte, rre = is_registration_successful(get_P_from_Rt(p_rot.as_matrix(), p_t),
                                                  get_P_from_Rt(g_rot.as_matrix(), g_t))
        if ok:
            ok_n, rte_sum, rre_sum = ok_n + 1, rte_sum + rte, rre_sum + rre
            if verbose:
                print(p_row)
    rate = ok_n / len(gt)
    avg_rte, avg_rre = rte_sum / ok_n, rre_sum / ok_n        # ZeroDivisionError without a success, as the reference
    print("Registration successful rate: %.2f, successful counter: %d, \n"
          "average Relative Translation Error (RTE): %.2f, average Relative Rotation Error (RRE): %.2f"
          % (rate, ok_n, avg_rte, avg_rre))
    return rate, avg_rte, avg_rre
