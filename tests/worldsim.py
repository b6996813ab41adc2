"""TEST INFRASTRUCTURE.  World-frame scenarios for the ORBmatcher methods whose host side projects
MapPoints itself (reference src/ORBmatcher.cc:311-427, 903-1207, 1212-1455, 1622-1759), and a numpy
float32 restatement of those host-side gates.  The restatement follows the reference expression by
expression (plain 3x3 * 3x1 cv::Mat products in float, transposed products and norms in double), so
that the projected view it produces can be handed to the oracle and compared with what the C++
drop-in class (include/ORBmatcher.h) computes from the same MapPoints."""
import numpy as np

F32, F64 = np.float32, np.float64
CAM = (718.856, 718.856, 607.19, 185.2, 386.1448)   # fx fy cx cy bf  (Examples/Stereo/KITTI00-02.yaml)


def _rodrigues(w):
    th = np.linalg.norm(w)
    if th < 1e-12:
        return np.eye(3)
    k = w / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)


def mul_row(R, x):
    """R * x on CV_32F data, 3x3 * 3x1 without transpose flags: cv::gemm's unrolled path, float products
    summed left to right in float (probed against cv2 4.13, tests/test_oracle_primitives.py)."""
    R = R.astype(F32); x = x.astype(F32)
    cols = [((R[r, 0] * x[:, 0] + R[r, 1] * x[:, 1]).astype(F32) + R[r, 2] * x[:, 2]).astype(F32) for r in range(3)]
    return np.stack(cols, 1)


def mul_row_add(R, x, t):
    """R * x + t: the product above, then one float add."""
    return (mul_row(R, x) + t.astype(F32)[None, :]).astype(F32)


def camera_centre(R, t):
    """-R.t() * t"""
    R = R.astype(F64); t = t.astype(F64)
    return np.array([-((R[0, r] * t[0] + R[1, r] * t[1]) + R[2, r] * t[2]) for r in range(3)]).astype(F32)


def norm3(v):
    v = v.astype(F64)
    return np.sqrt((v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1]) + v[:, 2] * v[:, 2]).astype(F32)


def decompose_sim3(S):
    sR, st = S[:3, :3].astype(F32), S[:3, 3].astype(F32)
    scw = F32(np.sqrt((F64(sR[0, 0]) * sR[0, 0] + F64(sR[0, 1]) * sR[0, 1]) + F64(sR[0, 2]) * sR[0, 2]))
    inv = F32(1.0 / F64(scw))
    return (sR * inv).astype(F32), (st * inv).astype(F32)


def predict_scale(maxd, dist, sf):
    log_sf = np.log(F32(sf[1]))
    ratio = (maxd.astype(F32) / dist.astype(F32)).astype(F32)
    n = np.ceil((np.log(ratio).astype(F32) / log_sf).astype(F32)).astype(np.int32)
    return np.clip(n, 0, len(sf) - 1)


def make_world(f, idx, n_pts, scale=1.0, match_frac=0.7, with_null=False, desc_src=None, perturb=None, src=None):
    """MapPoints in world coordinates that project (under the returned pose) onto / near keypoints of
    the frame arrays `f`; a few percent fail each of the reference's gates."""
    rng = np.random.default_rng(0x5EED0000 + 0x900000 + int(idx))
    fx, fy, cx, cy, bf = CAM
    n_kp = len(f["x"])
    R = _rodrigues(rng.normal(0, 0.05, 3)); t = rng.normal(0, 0.5, 3)
    T = np.eye(4); T[:3, :3] = scale * R; T[:3, 3] = scale * t
    T = T.astype(F32)
    rnd_src = rng.integers(0, n_kp, n_pts)
    src = rnd_src if src is None else np.asarray(src)
    matched = rng.random(n_pts) < match_frac
    u0 = np.where(matched, f["x"][src] + rng.normal(0, 2, n_pts), rng.uniform(-40, f["bounds"][2] + 40, n_pts))
    v0 = np.where(matched, f["y"][src] + rng.normal(0, 2, n_pts), rng.uniform(-40, f["bounds"][3] + 40, n_pts))
    z = rng.uniform(4, 40, n_pts) * np.where(rng.random(n_pts) < 0.03, -1.0, 1.0)
    Xc = np.stack([(u0 - cx) * z / fx, (v0 - cy) * z / fy, z], 1)
    Xw = ((Xc - t) @ R).astype(F32)                      # R^T (Xc - t)
    Ow = -R.T @ t
    PO = Xw.astype(F64) - Ow
    dist = np.linalg.norm(PO, axis=1)
    pred = np.clip(f["octave"][src] + rng.integers(0, 2, n_pts), 0, len(f["sf"]) - 1)
    maxd = dist * 1.2 ** (pred - 0.5) * np.where(rng.random(n_pts) < 0.05, 0.4, 1.0)
    mind = maxd / 1.2 ** 7
    normal = PO / dist[:, None] * np.where(rng.random(n_pts) < 0.08, -1.0, 1.0)[:, None]
    if desc_src is None:
        desc_src = f["desc"]
    desc = np.where(matched[:, None], perturb(rng, desc_src[src]) if perturb else desc_src[src],
                    rng.integers(0, 256, (n_pts, 32), dtype=np.uint8)).astype(np.uint8)
    dangle = np.where(rng.random(n_pts) < 0.8, rng.normal(30, 5, n_pts), rng.uniform(0, 360, n_pts))
    return dict(pose=T, xw=Xw, maxd=maxd.astype(F32), mind=mind.astype(F32), normal=normal.astype(F32), desc=desc,
                bad=rng.random(n_pts) < 0.04, null=(rng.random(n_pts) < 0.03) if with_null else np.zeros(n_pts, bool),
                nobs=rng.integers(1, 6, n_pts).astype(np.int32), found=rng.random(n_pts) < 0.04,
                angle=np.mod(f["angle"][src] + dangle, 360).astype(F32), src=src, R=R, t=t)


def blob(wd, prefix="p_"):
    u8 = lambda a: np.asarray(a).astype(np.uint8)
    return {prefix + "xw": wd["xw"].reshape(-1), prefix + "maxd": wd["maxd"], prefix + "mind": wd["mind"],
            prefix + "normal": wd["normal"].reshape(-1), prefix + "desc": wd["desc"].reshape(-1), prefix + "bad": u8(wd["bad"]),
            prefix + "null": u8(wd["null"]), prefix + "nobs": wd["nobs"], prefix + "found": u8(wd["found"]),
            prefix + "angle": wd["angle"]}


def _pts(valid, u, v, ur, radius, lo, hi, wd, with_angle):
    return dict(valid=valid, u=u.astype(F32), v=v.astype(F32), ur=None if ur is None else ur.astype(F32), radius=radius.astype(F32),
                min_level=lo.astype(np.int32), max_level=hi.astype(np.int32), desc=wd["desc"], angle=wd["angle"] if with_angle else None)


def gate_relocalisation(f, wd, th):
    """SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist), reference :1626-1672."""
    fx, fy, cx, cy = (F32(c) for c in CAM[:4])
    R, t = wd["pose"][:3, :3], wd["pose"][:3, 3]
    Ow = camera_centre(R, t)
    with np.errstate(all="ignore"):
        pc = mul_row_add(R, wd["xw"], t)
        invz = (1.0 / pc[:, 2].astype(F64)).astype(F32)
        u = (fx * pc[:, 0]) * invz + cx
        v = (fy * pc[:, 1]) * invz + cy
        valid = ~wd["null"] & ~wd["bad"] & ~wd["found"]
        b = f["bounds"]
        valid &= ~((u < F32(b[0])) | (u > F32(b[2]))) & ~((v < F32(b[1])) | (v > F32(b[3])))
        dist = norm3(wd["xw"] - Ow)
        valid &= ~((dist < F32(0.8) * wd["mind"]) | (dist > F32(1.2) * wd["maxd"]))
        lvl = predict_scale(wd["maxd"], dist, f["sf"])
    return _pts(valid, u, v, None, F32(th) * f["sf"][lvl], lvl - 1, lvl + 1, wd, True)


def gate_keyframe(f, wd, th, sim3, skip, bf=None):
    """Shared gate of SearchByProjection(KeyFrame*, Scw, ...) (:320-380), Fuse (:905-972) and Fuse(Scw) (:1068-1127):
    positive depth, IsInImage, distance range, viewing angle; levels [pred - 1, pred]."""
    fx, fy, cx, cy = (F32(c) for c in CAM[:4])
    if sim3:
        R, t = decompose_sim3(wd["pose"])
    else:
        R, t = wd["pose"][:3, :3], wd["pose"][:3, 3]
    Ow = camera_centre(R, t)
    with np.errstate(all="ignore"):
        pc = mul_row_add(R, wd["xw"], t)
        valid = ~skip & ~(pc[:, 2] < 0)
        invz = F32(1) / pc[:, 2]
        x = pc[:, 0] * invz; y = pc[:, 1] * invz
        u = fx * x + cx; v = fy * y + cy
        b = f["bounds"]
        valid &= (u >= F32(b[0])) & (u < F32(b[2])) & (v >= F32(b[1])) & (v < F32(b[3]))
        PO = wd["xw"] - Ow
        dist = norm3(PO)
        valid &= ~((dist < F32(0.8) * wd["mind"]) | (dist > F32(1.2) * wd["maxd"]))
        PO64, Pn = PO.astype(F64), wd["normal"].astype(F64)
        dot = (PO64[:, 0] * Pn[:, 0] + PO64[:, 1] * Pn[:, 1]) + PO64[:, 2] * Pn[:, 2]
        valid &= ~(dot < 0.5 * dist.astype(F64))
        lvl = predict_scale(wd["maxd"], dist, f["sf"])
        ur = None if bf is None else u - F32(bf) * invz
    return _pts(valid, u, v, ur, F32(th) * f["sf"][lvl], lvl - 1, lvl, wd, False)


def gate_sim3(dst, wd, Rw, tw, sR, tt, th, skip):
    """One direction of SearchBySim3 (reference :1259-1305 / :1337-1383): MapPoints seen in one keyframe
    (pose Rw, tw) moved by the similarity (sR, tt) into the other one (frame arrays `dst`)."""
    fx, fy, cx, cy = (F32(c) for c in CAM[:4])
    with np.errstate(all="ignore"):
        pa = mul_row_add(Rw, wd["xw"], tw)
        pb = mul_row_add(sR, pa, tt)
        valid = ~skip & ~(pb[:, 2] < 0)
        invz = (1.0 / pb[:, 2].astype(F64)).astype(F32)
        x = pb[:, 0] * invz; y = pb[:, 1] * invz
        u = fx * x + cx; v = fy * y + cy
        b = dst["bounds"]
        valid &= (u >= F32(b[0])) & (u < F32(b[2])) & (v >= F32(b[1])) & (v < F32(b[3]))
        dist = norm3(pb)
        valid &= ~((dist < F32(0.8) * wd["mind"]) | (dist > F32(1.2) * wd["maxd"]))
        lvl = predict_scale(wd["maxd"], dist, dst["sf"])
    return _pts(valid, u, v, None, F32(th) * dst["sf"][lvl], lvl - 1, lvl, wd, False)


def sim3_matrices(s12, R12, t12):
    """sR12, sR21, t21 of reference :1229-1231."""
    s12 = F32(s12); R12 = R12.astype(F32); t12 = t12.astype(F32)
    sR12 = (s12 * R12).astype(F32)
    sR21 = (F32(1.0 / F64(s12)) * R12.T).astype(F32)
    t21 = (-mul_row(sR21, t12[None, :])[0]).astype(F32)      # -sR21 * t12: plain product, float
    return sR12, sR21, t21


def frustum_scenario(idx, n_pts=3000):
    """Frame + world MapPoints for Frame::isInFrustum: the make_world points (a few percent fail each gate)."""
    import importlib
    synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
    f = synth.make_frame_arrays(2000, 560 + idx)
    wd = make_world(f, 40 + idx, n_pts, perturb=synth.perturb_descriptors)
    return f, wd


def distinctive_scenario(idx, n_kf=24, rows=300, n_points=400):
    """KeyFrames with descriptor rows and MapPoints observing 0..n_kf of them: each point's rows are noisy copies of
    one base descriptor (some points get coarse descriptors, so medians tie)."""
    rng = np.random.default_rng(0x5EED0000 + 0xD0000 + idx)
    kf_desc = rng.integers(0, 256, (n_kf, rows, 32), dtype=np.uint8)
    start = [0]; okf = []; orow = []
    free = [list(rng.permutation(rows)) for _ in range(n_kf)]
    for p in range(n_points):
        n_obs = int(rng.choice([0, 1, 2, 3, 5, 8, 13, n_kf], p=[0.02, 0.05, 0.1, 0.2, 0.25, 0.2, 0.13, 0.05]))
        kfs = rng.permutation(n_kf)[:n_obs]
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        coarse = rng.random() < 0.3
        for k in kfs:
            if not free[k]:
                continue
            r = free[k].pop()
            flips = np.packbits(rng.random(256) < 0.1)
            if coarse:
                flips = flips & 0x11
            kf_desc[k, r] = base ^ flips
            okf.append(int(k)); orow.append(int(r))
        start.append(len(okf))
    return kf_desc, np.array(start, np.int32), np.array(okf, np.int32), np.array(orow, np.int32)
