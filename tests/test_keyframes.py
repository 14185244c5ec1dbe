"""The keyframe-structured landmark query (ORBSLAM3Interface::mapPointsVisibleFromPose as the service runs it,
ORBW/src/orb_slam3_interface.cpp:188-317): per-keyframe prefilter, every map point once, and the maxLandmarks cutoff that also
applies in exhaustive mode (:270-271).  CPU: hand-derivable cases against the oracle's restatement.  GPU: the device query and
isPoseSafe behind it against the oracle on a synthetic atlas of 240 keyframes, with the cutoff active."""
import numpy as np
import pytest


def _pose(capi, x=0.0, y=0.0, z=0.3, yaw=0.0):
    p = np.zeros(1, dtype=capi.POSE_DTYPE)
    p["x"], p["y"], p["z"] = x, y, z
    p["qz"], p["qw"] = np.sin(yaw / 2), np.cos(yaw / 2)
    return p


def _kf(capi, x, y, yaw, first, count, z=0.3):
    perm = np.array([[0, -1, 0], [0, 0, -1], [1, 0, 0]], dtype=np.float64)
    c, s = np.cos(yaw), np.sin(yaw)
    R = np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])
    k = np.zeros(1, dtype=capi.KEYFRAME_DTYPE)
    k["Rwc"] = (perm @ R @ perm.T).astype(np.float32).ravel()
    k["twc"] = (perm @ np.array([x, y, z])).astype(np.float32)
    k["first"], k["count"] = first, count
    return k


def test_oracle_keyframe_query_known_answers(capi, oracle):
    # 12 points straight ahead of a camera at the origin looking along +x (all visible), 3 behind it (never visible)
    ahead = np.stack([np.linspace(2, 6, 12), np.zeros(12), np.full(12, 0.3)], axis=1)
    behind = np.array([[-2, 0, 0.3], [-3, 0.5, 0.3], [-4, -0.5, 0.3]])
    pts = np.concatenate([ahead, behind]).astype(np.float32)
    pose = _pose(capi)
    # three keyframes at the camera's own pose: points 0-4 | 3-8 (3, 4 seen before) + two behind | 9-11
    lists = [np.arange(0, 5), np.array([3, 4, 5, 6, 7, 8, 12, 13]), np.arange(9, 12)]
    kfs, at = [], 0
    for l in lists:
        kfs.append(_kf(capi, 0, 0, 0, at, len(l)))
        at += len(l)
    kfs = np.concatenate(kfs)
    idx = np.concatenate(lists).astype(np.uint32)
    # no cutoff in reach: every visible point once, in keyframe order then list order; pose frame = map frame minus (0, 0, 0.3)
    out, chk = oracle.landmarks_in_view_keyframes(pts, kfs, idx, pose, max_landmarks=1000, exhaustive=True)
    assert chk == 3 and len(out) == 12
    np.testing.assert_allclose(out[:, 0], np.linspace(2, 6, 12), rtol=1e-6)
    np.testing.assert_allclose(out[:, 2], 0.0, atol=1e-6)
    # the cutoff is "MORE than maxLandmarks collected" at the start of a keyframe, in exhaustive mode too (:270-271):
    # after keyframe 0: 5 collected; 5 > 4 stops before keyframe 1, which still counts as checked
    out, chk = oracle.landmarks_in_view_keyframes(pts, kfs, idx, pose, max_landmarks=4, exhaustive=True)
    assert len(out) == 5 and chk == 2
    # 5 > 5 is false: keyframe 1 runs whole (a keyframe is never cut in the middle): 5 + 4 fresh visible ones; 9 > 5 stops before 2
    out, chk = oracle.landmarks_in_view_keyframes(pts, kfs, idx, pose, max_landmarks=5, exhaustive=True)
    assert len(out) == 9 and chk == 3
    # prefilter (exhaustive_search = false), max_distance 5: a keyframe 12 m away (>= 2 * max_distance) is skipped and is not
    # counted as checked; one at 7 m behind the camera (5 < d < 10, z < 0) is skipped; one at 7 m ahead facing the same way
    # (relative yaw 0 lies in [0, pi/2)) is skipped too; one at 7 m ahead facing back at the camera (relative yaw pi) is kept
    far = _kf(capi, 12, 0, 0, 0, 5)
    back = _kf(capi, -7, 0, 0, 0, 5)
    ahead_same = _kf(capi, 7, 0, 0, 0, 5)
    ahead_facing = _kf(capi, 7, 0, np.pi, 0, 5)
    for k, want_checked in ((far, 0), (back, 0), (ahead_same, 0), (ahead_facing, 1)):
        out, chk = oracle.landmarks_in_view_keyframes(pts, k, idx[:5], pose, max_distance=5.0, exhaustive=False)
        assert chk == want_checked and len(out) == 5 * want_checked, (k["twc"], chk, len(out))
    # within max_distance and behind the camera: skipped only when the relative yaw is in [2.0944, 4.18879)
    near_behind_facing_away = _kf(capi, -2, 0, np.pi, 0, 5)
    near_behind_same = _kf(capi, -2, 0, 0, 0, 5)
    out, chk = oracle.landmarks_in_view_keyframes(pts, near_behind_facing_away, idx[:5], pose, max_distance=5.0, exhaustive=False)
    assert chk == 0
    out, chk = oracle.landmarks_in_view_keyframes(pts, near_behind_same, idx[:5], pose, max_distance=5.0, exhaustive=False)
    assert chk == 1 and len(out) == 5
    # one keyframe holding every point = the query the other entry points answer
    one = _kf(capi, 0, 0, 0, 0, len(pts))
    out, chk = oracle.landmarks_in_view_keyframes(pts, one, np.arange(len(pts), dtype=np.uint32), pose, exhaustive=True)
    want = oracle.landmarks_in_view(pts, pose)
    assert np.array_equal(out, want)


@pytest.mark.gpu
@pytest.mark.parametrize("exhaustive", [True, False])
def test_gpu_keyframe_query_matches_oracle(capi, oracle, synth, exhaustive):
    rng = np.random.default_rng(11)
    pts = rng.uniform(-25, 25, size=(120_000, 3)).astype(np.float32)
    pts[:, 2] = rng.uniform(0, 3, len(pts))
    kfs, idx = synth.make_keyframes(pts, 240, seed=5)
    assert len(kfs) == 240 and len(idx) > 100_000
    oracle.fisher_generate()
    ctx = capi.Context()
    try:
        ctx.fisher_generate_lookup()
        ctx.set_map_points(pts)
        with pytest.raises(capi.FitslamError):  # a list that points outside the resident points
            ctx.set_keyframes(kfs[:1], np.array([len(pts)] * int(kfs[0]["count"]), np.uint32))
        ctx.set_keyframes(kfs, idx)
        cut = 0
        for t in range(24):
            k = int(rng.integers(len(kfs)))  # query poses near keyframes: something is in view
            perm = np.array([[0, -1, 0], [0, 0, -1], [1, 0, 0]], dtype=np.float64)
            t_ros = perm.T @ kfs[k]["twc"].astype(np.float64)
            pose = _pose(capi, t_ros[0] + rng.normal(0, 1.0), t_ros[1] + rng.normal(0, 1.0), 0.3, rng.uniform(-np.pi, np.pi))
            for max_landmarks in (1000, 150, 0):
                want, chk_w = oracle.landmarks_in_view_keyframes(pts, kfs, idx, pose, max_landmarks, 5.0, exhaustive)
                st, got, n, chk = ctx.landmarks_in_view_keyframes(pose, max_landmarks, 5.0, exhaustive)
                assert st == 0 and n == len(want) and chk == chk_w, (t, max_landmarks, n, len(want), chk, chk_w)
                assert np.array_equal(got, want, equal_nan=True)
                # the cutoff really fired where expected: fewer landmarks than the one-keyframe query sees
                if max_landmarks == 1000 and n > 1000:
                    cut += 1
                    full = oracle.landmarks_in_view(pts, pose)
                    assert n <= len(full)
                # isPoseSafe behind the service: same landmarks, same voxels; information within 1e-4 of the reference's float sum
                sto, winfo, winfo64, wsafe, wn, wused = oracle.fisher_pose_keyframes(pts, kfs, idx, pose, max_landmarks, 5.0, exhaustive)
                info, safe, n2, used = ctx.score_fisher_keyframes(pose, max_landmarks, 5.0, exhaustive)
                assert sto == 0 and n2 == wn == n and used == wused
                assert abs(info - winfo64) <= 1e-6 * abs(winfo64) + 1e-6
                assert abs(info - winfo) <= 1e-4 * abs(winfo) + 1e-6
                if abs(winfo - 550.0) > 1.0:
                    assert safe == wsafe
            if n and t == 0:  # a short buffer: the count is whole, the first entries are written
                st, part, n3, _ = ctx.landmarks_in_view_keyframes(pose, 0, 5.0, exhaustive, capacity=1)
                assert (st == capi.E_CAPACITY) == (n3 > 1) and np.array_equal(part[:1], want[:1], equal_nan=True)
        assert not exhaustive or cut > 0  # the 1000-landmark cutoff was active in the exhaustive run
        # new map points drop the atlas (its indices meant the old points)
        ctx.set_map_points(pts[:100])
        with pytest.raises(capi.FitslamError):
            ctx.landmarks_in_view_keyframes(pose)
    finally:
        ctx.close()
