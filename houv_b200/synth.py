"""Synthetic MVP-shaped registration pairs (the real MVP_*_RG.h5 files are not
shipped with the reference: registration/data/.gitignore).

Pose sampling restates the reference's generator
(/root/reference/registration/dataset.py:16-37 random_pose / random_rotation /
random_translation, and :297-301 how MVP_RG composes them):
    transform = random_pose(max_angle, max_trans / 2)
    pose1     = random_pose(pi, max_trans);  pose2 = transform @ pose1
    src = partial1 @ pose1[:3,:3].T + pose1[:3,3];  tgt = partial2 @ pose2[:3,:3].T + pose2[:3,3]
with cfgs/houv.yaml:20-21 max_angle = 180 deg, max_trans = 0.5, and ground truth
`transform` (src -> tgt).  Shapes: two different 2048-point partial views of
one closed surface inside the radius-0.5 ball, like MVP.  80 % of the pairs
draw the rotation angle from [0, 45] deg and 20 % from [0, 180] deg
(registration/README.md:56-57).
"""
import numpy as np


def random_rotation(rng, max_angle):
    axis = rng.randn(3)
    axis /= np.linalg.norm(axis)
    angle = rng.rand() * max_angle
    A = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    R = np.eye(3) + np.sin(angle) * A + (1 - np.cos(angle)) * np.dot(A, A)
    return R, angle


def random_translation(rng, max_dist):
    t = rng.randn(3)
    t /= np.linalg.norm(t)
    t *= rng.rand() * max_dist
    return np.expand_dims(t, 1)


def random_pose(rng, max_angle, max_trans):
    R, angle = random_rotation(rng, max_angle)
    t = random_translation(rng, max_trans)
    return np.concatenate([np.concatenate([R, t], 1), [[0, 0, 0, 1]]], 0), angle


def _surface(rng, n):
    """n points on the union of three random ellipsoids, scaled into the 0.5-ball."""
    pts = []
    for _ in range(3):
        c = rng.uniform(-0.18, 0.18, 3)
        r = rng.uniform(0.08, 0.3, 3)
        Rm, _ = random_rotation(rng, np.pi)
        u = rng.randn(n // 3 + 1, 3)
        u /= np.linalg.norm(u, axis=1, keepdims=True)
        pts.append((u * r) @ Rm.T + c)
    p = np.concatenate(pts)[:n]
    p -= p.mean(0)
    p *= 0.5 / np.abs(np.linalg.norm(p, axis=1)).max()
    return p


def _box_surface(rng, n, size, centre):
    """n points uniform on the surface of an axis-aligned box."""
    size = np.asarray(size, float)
    areas = np.array([size[1] * size[2], size[0] * size[2], size[0] * size[1]])
    axis = rng.choice(3, n, p=areas / areas.sum())
    pts = (rng.rand(n, 3) - 0.5) * size
    pts[np.arange(n), axis] = np.where(rng.rand(n) < 0.5, -0.5, 0.5) * size[axis]
    return pts + np.asarray(centre, float)


def _chair(rng, n):
    """n points on a chair-like object (seat, back, four legs, ONE armrest): no rotational or
    mirror symmetry, so a pose in the wrong basin leaves a large trimmed Chamfer distance --
    the case HOUV's 0.030 retry threshold is there for (houv.py:154-157).  Proportions are drawn
    per object; scaled into the 0.5-ball like MVP."""
    w, d, t = rng.uniform(0.5, 0.8), rng.uniform(0.5, 0.8), 0.05
    h, hl, r = rng.uniform(0.5, 0.9), rng.uniform(0.35, 0.6), 0.05
    parts = [((w, d, t), (0, 0, 0)),                                  # seat
             ((w, t, h), (0, -d / 2 + t / 2, h / 2)),                 # back
             ((t, d * 0.8, t), (w / 2 - t / 2, 0, rng.uniform(0.2, 0.3)))]  # one armrest
    for sx in (-1, 1):
        for sy in (-1, 1):
            parts.append(((r, r, hl), (sx * (w / 2 - r), sy * (d / 2 - r), -hl / 2)))
    area = np.array([2 * (a * b + b * c + a * c) for (a, b, c), _ in parts])
    cnt = rng.multinomial(n, area / area.sum())
    p = np.concatenate([_box_surface(rng, k, sz, c) for k, (sz, c) in zip(cnt, parts)])
    Rm, _ = random_rotation(rng, np.pi)
    p = p @ Rm.T
    p -= p.mean(0)
    p *= 0.5 / np.linalg.norm(p, axis=1).max()
    return p


def _partial(rng, full, n):
    """the n points facing a random view direction most (a crude visible side), shuffled"""
    d = rng.randn(3)
    d /= np.linalg.norm(d)
    score = full @ d + 0.05 * rng.randn(len(full))
    keep = np.argsort(-score)[: int(len(full) * 0.6)]
    sel = rng.choice(keep, n, replace=False)
    return full[sel]


def make_pair(p, n_points=2048, max_angle=np.pi, max_trans=0.5, seed=2021, shape="ellipsoids",
              radius=0.5):
    """-> src[n,3], tgt[n,3], transform[4,4] (float32), deterministic in p.
    shape: "ellipsoids" (three random ellipsoids: smooth, near-symmetric -- a wrong-basin pose
    often stays BELOW the 0.030 retry threshold) or "chair" (asymmetric, see _chair)."""
    rng = np.random.RandomState(seed + p)
    full = _chair(rng, 8 * n_points) if shape == "chair" else _surface(rng, 8 * n_points)
    if radius != 0.5:  # object size only; the pose distribution (translations) stays MVP's
        full = full * (radius / 0.5)
    part1 = _partial(rng, full, n_points)
    part2 = _partial(rng, full, n_points)
    ang_cap = max_angle if rng.rand() < 0.2 else min(max_angle, np.pi / 4)
    transform, _ = random_pose(rng, ang_cap, max_trans / 2)
    pose1, _ = random_pose(rng, np.pi, max_trans)
    pose2 = transform @ pose1
    src = part1 @ pose1[:3, :3].T + pose1[:3, 3]
    tgt = part2 @ pose2[:3, :3].T + pose2[:3, 3]
    f = np.float32
    return src.astype(f), tgt.astype(f), transform.astype(f)


def make_pairs(first, count, n_points=2048, **kw):
    """Pairs [first, first+count): src[count,n,3], tgt[count,n,3], transform[count,4,4]."""
    out = [make_pair(first + i, n_points, **kw) for i in range(count)]
    return tuple(np.stack([o[k] for o in out]) for k in range(3))


def make_overlap_pair(p, n_points=2048, max_angle=np.pi / 4.5, max_trans=0.5, seed=7000):
    """A well-conditioned pair for end-state parity checks (SURVEY §7 hard part 1 (iii)): both
    clouds are different random subsets of the SAME partial view (full overlap, no exact
    correspondences), relative rotation below `max_angle` (40 deg: inside what angle_base 0 can
    reach, houv.py:96).  -> src[n,3], tgt[n,3], transform[4,4] (float32), deterministic in p."""
    rng = np.random.RandomState(seed + p)
    full = _surface(rng, 8 * n_points)
    d = rng.randn(3)
    d /= np.linalg.norm(d)
    keep = np.argsort(-(full @ d))[: int(len(full) * 0.6)]
    part1 = full[rng.choice(keep, n_points, replace=False)]
    part2 = full[rng.choice(keep, n_points, replace=False)]
    transform, _ = random_pose(rng, max_angle, max_trans / 2)
    pose1, _ = random_pose(rng, np.pi, max_trans)
    pose2 = transform @ pose1
    src = part1 @ pose1[:3, :3].T + pose1[:3, 3]
    tgt = part2 @ pose2[:3, :3].T + pose2[:3, 3]
    f = np.float32
    return src.astype(f), tgt.astype(f), transform.astype(f)
