"""CPU: the distCUDA2 restatement (oracle/knn_ref.c) against known answers and an independent numpy computation."""
import numpy as np

from oracle import knn_oracle


def test_known_answers():
    # unit-spaced points on a line: interior points see 1, 1, 4; the ends see 1, 4, 9
    pts = np.zeros((6, 3), np.float32)
    pts[:, 0] = np.arange(6)
    out = knn_oracle.mean_dist2(pts)
    assert out[0] == np.float32(14.0) / np.float32(3.0) and out[5] == out[0]
    assert out[2] == np.float32(6.0) / np.float32(3.0) and out[1] == np.float32(6.0) / np.float32(3.0)


def test_duplicates_count_with_distance_zero_and_self_is_excluded():
    pts = np.array([[0, 0, 0], [0, 0, 0], [1, 0, 0], [0, 2, 0], [5, 5, 5]], np.float32)
    out = knn_oracle.mean_dist2(pts)
    assert out[0] == np.float32(5.0) / np.float32(3.0)      # 0 (the duplicate), 1, 4
    assert out[0] == out[1]


def test_fewer_than_four_points_keep_flt_max_like_the_reference():
    out = knn_oracle.mean_dist2(np.array([[0, 0, 0], [1, 0, 0]], np.float32))
    assert np.isinf(out).all() or (out > 1e37).all()
    assert knn_oracle.mean_dist2(np.zeros((0, 3), np.float32)).shape == (0,)


def test_matches_numpy_float64_three_smallest():
    rng = np.random.default_rng(0)
    pts = rng.uniform(-1.3, 1.3, (700, 3)).astype(np.float32)
    d2 = ((pts[:, None, :].astype(np.float64) - pts[None, :, :]) ** 2).sum(-1)
    np.fill_diagonal(d2, np.inf)
    want = np.sort(d2, axis=1)[:, :3].sum(1) / 3.0
    got = knn_oracle.mean_dist2(pts)
    assert np.abs(got - want).max() <= 1e-6 * want.max()
