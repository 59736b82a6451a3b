"""Synthetic RGB-D scenes for the benchmark configurations of SURVEY.md section 8(d) (row S1): an analytic room of
oriented boxes and ellipsoids, ray-cast depth / RGB / instance-id frames, per-object memory clouds and DINO-like
embeddings.  Replaces the reference's habitat-sim / dataset inputs (hm3d_data_generation.py; GL rendering is
unavailable on this box).  Plain tensor compute (torch on CPU or CUDA); it only PRODUCES inputs and is never timed.

Conventions follow the reference: a pixel (row r, col c) back-projects to X = linspace(-W/2, W/2, W)[c] * d / fx,
Y = linspace(H/2, -H/2, H)[r] * d / fy, Z = d (utils/depth_utils.py:57-67), depth is float32 metres along +Z of the
camera frame; the pose that localise() recovers maps camera-frame points into the memory (world) frame.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np
import torch


@dataclass
class Scene:
    """An object is a union of primitives (oriented boxes / ellipsoids): a base standing on the floor plus parts attached
    to its surface at random 3-d orientations, so that objects have edges, corners and junctions (FPFH has something to
    match) and no rotational symmetry (a pose is identifiable).  Arrays are per PRIMITIVE; ``obj`` maps a primitive to
    its object."""
    kind: np.ndarray          # [n_prim] 0 = box, 1 = ellipsoid
    half: np.ndarray          # [n_prim,3] half extents / semi axes
    R: np.ndarray             # [n_prim,3,3] primitive -> world
    t: np.ndarray             # [n_prim,3]
    albedo: np.ndarray        # [n_prim,3]
    room: float               # half side of the square room
    height: float             # ceiling height
    emb_centres: np.ndarray   # [m,D]
    memory_embs: list = field(default_factory=list)   # list of [n_e, D] float32
    seed: int = 0
    obj: np.ndarray = None    # [n_prim] object of each primitive
    centre: np.ndarray = None  # [m,3] object centres (base primitive)

    def __post_init__(self):
        if self.obj is None:
            self.obj = np.arange(len(self.kind))
        if self.centre is None:
            self.centre = np.array([self.t[self.obj == j][0] for j in range(self.m)])

    @property
    def m(self):
        return int(self.obj.max()) + 1 if len(self.obj) else 0

    @property
    def n_prim(self):
        return len(self.kind)

    def prims_of(self, j):
        return np.nonzero(self.obj == j)[0]


LIGHT = np.array([0.35, 0.45, 0.82]) / np.linalg.norm([0.35, 0.45, 0.82])


def _yaw(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1.0]])


def _random_rotation(rng):
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def _prim_min_z(kind, half, R, t):
    """lowest world z of a primitive."""
    if kind == 0:
        return t[2] - np.abs(R[2]) @ half
    return t[2] - np.sqrt(((R[2] * half) ** 2).sum())


def make_scene(m=40, seed=1002, D=768, size=(0.3, 0.9), n_emb=(5, 20), sigma=0.0175, common=0.65, parts=(2, 4)):
    """m objects on a jittered floor grid inside a square room (world frame: z up).  ``parts`` = (min, max) number of
    primitives attached to every object's base; (0, 0) gives the single-primitive objects of round 1."""
    rng = np.random.default_rng(seed)
    g = int(np.ceil(np.sqrt(m)))
    pitch = 1.9
    room = g * pitch / 2 + 1.0
    cells = rng.permutation(g * g)[:m]
    kind = (rng.random(m) < 0.4).astype(np.int64)
    half = rng.uniform(size[0] / 2, size[1] / 2, (m, 3))
    R = np.stack([_yaw(a) for a in rng.uniform(0, 2 * np.pi, m)])
    t = np.zeros((m, 3))
    t[:, 0] = (cells % g - (g - 1) / 2) * pitch + rng.uniform(-0.25, 0.25, m)
    t[:, 1] = (cells // g - (g - 1) / 2) * pitch + rng.uniform(-0.25, 0.25, m)
    t[:, 2] = half[:, 2]
    albedo = rng.uniform(0.25, 0.95, (m, 3))
    obj = np.arange(m)
    centre = t.copy()
    if parts[1] > 0:
        prng = np.random.default_rng(seed * 9176 + 11)        # own stream: the base layout above is unchanged by `parts`
        ek, eh, eR, et, ea, eo = [], [], [], [], [], []
        for j in range(m):
            for _ in range(int(prng.integers(parts[0], parts[1] + 1))):
                k = int(prng.random() < 0.35)
                h = half[j].mean() * prng.uniform(0.3, 0.75, 3)
                Rp = _random_rotation(prng)
                u = prng.normal(size=3)                        # anchor on the base's surface, upper half preferred
                u[2] = abs(u[2]) * 0.8
                u /= np.linalg.norm(u)
                if kind[j] == 0:
                    a = u / np.abs(u / half[j]).max()           # point of the box surface along u
                else:
                    a = u / np.linalg.norm(u / half[j])
                tp = t[j] + R[j] @ a
                lift = 0.01 - _prim_min_z(k, h, Rp, tp)
                if lift > 0:
                    tp = tp + np.array([0, 0, lift])
                ek.append(k); eh.append(h); eR.append(Rp); et.append(tp)
                ea.append(np.clip(albedo[j] + prng.uniform(-0.22, 0.22, 3), 0.15, 0.98)); eo.append(j)
        if ek:
            kind = np.concatenate([kind, np.array(ek, dtype=np.int64)])
            half = np.concatenate([half, np.array(eh)])
            R = np.concatenate([R, np.array(eR)])
            t = np.concatenate([t, np.array(et)])
            albedo = np.concatenate([albedo, np.array(ea)])
            obj = np.concatenate([obj, np.array(eo)])
    gdir = rng.normal(size=D)
    gdir /= np.linalg.norm(gdir)
    cent = rng.normal(size=(m, D))
    cent /= np.linalg.norm(cent, axis=1, keepdims=True)
    cent = common * gdir[None] + np.sqrt(1 - common ** 2) * cent
    cent /= np.linalg.norm(cent, axis=1, keepdims=True)
    mem = []
    for j in range(m):
        ne = int(rng.integers(n_emb[0], n_emb[1] + 1))
        e = cent[j][None] + rng.normal(0, sigma, (ne, D))
        mem.append((e * rng.uniform(5, 40, (ne, 1))).astype(np.float32))     # DINOv2 outputs are not normalised
    return Scene(kind, half, R, t, albedo, room, 2.8, cent, mem, seed, obj, centre)


def _texture(p):
    """smooth world-space pattern in [0.55, 1]: gives coloured ICP a photometric gradient."""
    s = torch.sin(9.0 * p[..., 0] + 1.3) * torch.cos(7.0 * p[..., 1] - 0.4) * torch.sin(8.0 * p[..., 2] + 0.7)
    return 0.775 + 0.225 * s


def _shade(albedo, normal_w, p_w):
    lam = (normal_w * torch.as_tensor(LIGHT, dtype=p_w.dtype, device=p_w.device)).sum(-1).clamp(min=0.0)
    return (albedo * (0.35 + 0.65 * lam)[..., None] * _texture(p_w)[..., None]).clamp(0.0, 1.0)


def camera_pose(scene, rng, n_try=200):
    """Random camera inside the room looking at an object from 1.6-3.2 m.  Returns T_wc (camera -> world) with the
    reference's back-projection axes: +X along image columns, +Y up (decreasing row), +Z forward (depth); det = +1."""
    for _ in range(n_try):
        j = int(rng.integers(0, scene.m))
        ang = rng.uniform(0, 2 * np.pi)
        dist = rng.uniform(1.6, 3.2)
        eye = scene.centre[j] + np.array([np.cos(ang) * dist, np.sin(ang) * dist, 0.0])
        eye[2] = rng.uniform(1.1, 1.6)
        if np.abs(eye[:2]).max() > scene.room - 0.3:
            continue
        d2 = np.linalg.norm(scene.centre[:, :2] - eye[None, :2], axis=1)
        if d2.min() < 0.9:
            continue
        tgt = scene.centre[j] + rng.uniform(-0.15, 0.15, 3)
        fwd = tgt - eye
        fwd /= np.linalg.norm(fwd)
        # The reference's back-projection (X grows with the column, Y with decreasing row, Z = depth) reads pixels into a
        # frame (x, y, z); for the recovered pose to be a rotation, (x, y, z) must be RIGHT-handed: x = up x fwd
        # (image columns run towards the camera's left), y = fwd x x.  Round 1 used x = fwd x up, a reflection: every
        # detected cloud was the mirror image of its memory cloud and no rigid registration could match the true pose.
        xc = np.cross(np.array([0, 0, 1.0]), fwd)
        xc /= np.linalg.norm(xc)
        yc = np.cross(fwd, xc)
        T = np.eye(4)
        T[:3, 0], T[:3, 1], T[:3, 2], T[:3, 3] = xc, yc, fwd, eye
        return T
    raise RuntimeError("could not place a camera")


def render(scene, T_wc, H=600, W=600, f=300.0, device="cpu", depth_noise=0.002, seed=0, use_kernel=False, want_geometry=False):
    """Ray cast one frame -> (depth float32 [H,W] metres, rgb uint8 [H,W,3], inst int32 [H,W] (-1 = background)).
    ``use_kernel`` (CUDA only) computes the first hits with the ibl_raycast_scene kernel instead of tensor ops; the two
    agree to rounding.  ``want_geometry`` returns the noise-free (hit depth f64 [P], inst [P], world normal [P,3]) instead."""
    dev = torch.device(device)
    dt = torch.float64
    hor = torch.as_tensor(np.linspace(-W / 2, W / 2, W, dtype=np.float32), dtype=dt, device=dev)
    ver = torch.as_tensor(np.linspace(H / 2, -H / 2, H, dtype=np.float32), dtype=dt, device=dev)
    dirs_c = torch.stack([hor[None, :].expand(H, W) / f, ver[:, None].expand(H, W) / f, torch.ones((H, W), dtype=dt, device=dev)], -1)
    Twc = torch.as_tensor(T_wc, dtype=dt, device=dev)
    o_w = Twc[:3, 3]
    d_w = dirs_c.reshape(-1, 3) @ Twc[:3, :3].T                    # NOT normalised: parameter == camera depth Z
    P = d_w.shape[0]
    if use_kernel:
        # the same first-hit computation as ONE CUDA kernel (csrc/raycast.cu, ibl_raycast_scene): thread per pixel
        from . import kernels as K
        with torch.cuda.device(dev):
            best, inst, nrm = K.raycast_scene(T_wc, H, W, f, scene.kind, scene.half, scene.R, scene.t, scene.room, scene.height)
    else:
        Ro = torch.as_tensor(scene.R, dtype=dt, device=dev)            # [m,3,3]
        to = torch.as_tensor(scene.t, dtype=dt, device=dev)
        hf = torch.as_tensor(scene.half, dtype=dt, device=dev)
        best = torch.full((P,), float("inf"), dtype=dt, device=dev)
        inst = torch.full((P,), -1, dtype=torch.int32, device=dev)
        nrm = torch.zeros((P, 3), dtype=dt, device=dev)
        # room: floor z=0, ceiling, 4 walls
        planes = [((0, 0, 1.0), 0.0), ((0, 0, -1.0), -scene.height), ((1.0, 0, 0), -scene.room), ((-1.0, 0, 0), -scene.room),
                  ((0, 1.0, 0), -scene.room), ((0, -1.0, 0), -scene.room)]
        for n_, off in planes:
            n_t = torch.as_tensor(n_, dtype=dt, device=dev)
            den = d_w @ n_t
            tt = (off - o_w @ n_t) / den
            ok = (den < 0) & (tt > 1e-6) & (tt < best)
            best = torch.where(ok, tt, best)
            nrm = torch.where(ok[:, None], n_t.expand(P, 3), nrm)
        for j in range(scene.n_prim):
            ol = (o_w - to[j]) @ Ro[j]                                  # world -> object: R^T (x - t)
            dl = d_w @ Ro[j]
            if scene.kind[j] == 0:                                      # box: slab method
                inv = 1.0 / dl
                t0 = (-hf[j] - ol) * inv
                t1 = (hf[j] - ol) * inv
                tn, tf = torch.minimum(t0, t1), torch.maximum(t0, t1)
                tnear, ax = tn.max(dim=1)
                tfar = tf.min(dim=1).values
                ok = (tnear < tfar) & (tnear > 1e-6) & (tnear < best)
                nl = torch.zeros((P, 3), dtype=dt, device=dev)
                nl.scatter_(1, ax[:, None], -torch.sign(dl.gather(1, ax[:, None])))
                tt = tnear
            else:                                                       # ellipsoid: unit sphere in scaled coordinates
                os_, ds_ = ol / hf[j], dl / hf[j]
                a = (ds_ * ds_).sum(1)
                b = (os_[None] * ds_).sum(1)
                c = (os_ * os_).sum() - 1.0
                disc = b * b - a * c
                tt = (-b - torch.sqrt(disc.clamp(min=0))) / a
                ok = (disc > 0) & (tt > 1e-6) & (tt < best)
                hit = ol + tt[:, None] * dl
                nl = hit / (hf[j] * hf[j])
                nl = nl / nl.norm(dim=1, keepdim=True).clamp(min=1e-12)
            best = torch.where(ok, tt, best)
            inst = torch.where(ok, torch.full_like(inst, j), inst)
            nrm = torch.where(ok[:, None], nl @ Ro[j].T, nrm)
    if want_geometry:
        return best, inst, nrm
    hit_w = o_w[None] + best[:, None] * d_w
    alb_tab = torch.as_tensor(np.concatenate([scene.albedo, [[0.62, 0.60, 0.57]]]), dtype=dt, device=dev)
    alb = alb_tab[inst.long()]                                      # -1 -> last row (room surfaces); per primitive
    obj_tab = torch.as_tensor(np.concatenate([scene.obj, [-1]]), dtype=torch.int32, device=dev)
    inst = obj_tab[inst.long()]                                     # primitive -> object id
    col = _shade(alb, nrm, hit_w)
    gen = torch.Generator(device="cpu").manual_seed(int(seed))
    col = (col + 0.01 * torch.randn((P, 3), generator=gen, dtype=dt).to(dev)).clamp(0, 1)
    depth = best + depth_noise * torch.randn((P,), generator=gen, dtype=dt).to(dev)
    depth = torch.where(torch.isfinite(best), depth, torch.zeros_like(depth)).clamp(min=0.0)
    rgb = (col * 255.0).round().to(torch.uint8).reshape(H, W, 3)
    return depth.to(torch.float32).reshape(H, W), rgb, inst.reshape(H, W)


def instance_masks(inst, max_objects=12, min_pixels=300):
    """Ground-truth instance masks eroded by 1 px, largest first -> (bool [n,1,H,W], object ids [n])."""
    H, W = inst.shape
    ids, counts = torch.unique(inst[inst >= 0], return_counts=True)
    order = torch.argsort(counts, descending=True)
    masks, keep_ids = [], []
    for k in order.tolist():
        if len(masks) >= max_objects:
            break
        mk = inst == ids[k]
        er = mk.clone()
        er[1:] &= mk[:-1]
        er[:-1] &= mk[1:]
        er[:, 1:] &= mk[:, :-1]
        er[:, :-1] &= mk[:, 1:]
        er[0] = er[-1] = False
        er[:, 0] = er[:, -1] = False
        if int(er.sum()) >= min_pixels:
            masks.append(er)
            keep_ids.append(int(ids[k]))
    if not masks:
        return torch.zeros((0, 1, H, W), dtype=torch.bool, device=inst.device), []
    return torch.stack(masks)[:, None], keep_ids


def query_embeddings(scene, obj_ids, rng, sigma=0.0175):
    D = scene.emb_centres.shape[1]
    e = scene.emb_centres[obj_ids] + rng.normal(0, sigma, (len(obj_ids), D))
    return (e * rng.uniform(5, 40, (len(obj_ids), 1))).astype(np.float32)


def _prim_surface(kind, h, n, rng, skip_bottom):
    """n random surface samples of a primitive in its own frame -> (points, outward normals)."""
    if kind == 0:
        areas = 4 * np.array([h[1] * h[2], h[1] * h[2], h[0] * h[2], h[0] * h[2], h[0] * h[1], h[0] * h[1]])
        if skip_bottom:
            areas = areas[:5]
        nf = len(areas)
        face = rng.choice(nf, n, p=areas / areas.sum())
        pl = rng.uniform(-1, 1, (n, 3)) * h
        nl = np.zeros((n, 3))
        ax = np.array([0, 0, 1, 1, 2, 2])[face]
        sg = np.array([1, -1, 1, -1, 1.0, -1.0])[face]
        pl[np.arange(n), ax] = sg * h[ax]
        nl[np.arange(n), ax] = sg
        return pl, nl
    u = rng.normal(size=(n, 3))
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    pl = u * h
    nl = pl / (h * h)
    nl /= np.linalg.norm(nl, axis=1, keepdims=True)
    return pl, nl


def _prim_area(kind, h, skip_bottom):
    if kind == 0:
        a = 8 * (h[1] * h[2] + h[0] * h[2] + h[0] * h[1])
        return a - (4 * h[0] * h[1] if skip_bottom else 0.0)
    return 4 * np.pi * ((((h[0] * h[1]) ** 1.6 + (h[0] * h[2]) ** 1.6 + (h[1] * h[2]) ** 1.6) / 3) ** (1 / 1.6))


def _inside(kind, h, R, t, pw, margin=1e-9):
    pl = (pw - t) @ R
    if kind == 0:
        return (np.abs(pl) < h - margin).all(axis=1)
    return ((pl / h) ** 2).sum(axis=1) < 1.0 - margin


def memory_cloud(scene, j, spacing=0.01, seed=0):
    """World-frame surface samples of object j (the visible surface of the union of its primitives: samples inside
    another primitive of the object or under the floor are dropped) with the same shading model as render()
    -> (pts f64, cols f64)."""
    rng = np.random.default_rng(seed * 100003 + j)
    prims = scene.prims_of(j)
    P, C = [], []
    for a, q in enumerate(prims):
        h = scene.half[q]
        skip_bottom = bool(a == 0 and scene.kind[q] == 0)         # the base box stands on the floor
        n = int(_prim_area(scene.kind[q], h, skip_bottom) / spacing ** 2)
        pl, nl = _prim_surface(scene.kind[q], h, n, rng, skip_bottom)
        pw = pl @ scene.R[q].T + scene.t[q]
        nw = nl @ scene.R[q].T
        keep = pw[:, 2] > 0.0
        for q2 in prims:
            if q2 != q:
                keep &= ~_inside(scene.kind[q2], scene.half[q2], scene.R[q2], scene.t[q2], pw)
        pw, nw = pw[keep], nw[keep]
        pw = pw + rng.normal(0, 0.001, pw.shape)
        col = _shade(torch.as_tensor(scene.albedo[q]), torch.as_tensor(nw), torch.as_tensor(pw)).numpy()
        col = np.clip(col + rng.normal(0, 0.01, col.shape), 0, 1)
        P.append(pw)
        C.append(col)
    return np.ascontiguousarray(np.concatenate(P)), np.ascontiguousarray(np.concatenate(C))


def make_frames(scene, n_frames, H=600, W=600, f=300.0, device="cpu", seed=0, max_objects=12, min_pixels=300):
    """-> list of dicts {rgb uint8 [H,W,3], depth f32 [H,W], masks bool [n,1,H,W], embs f32 [n,D], ids, T_wc} (host)."""
    rng = np.random.default_rng(scene.seed * 7919 + seed)
    frames = []
    while len(frames) < n_frames:
        T = camera_pose(scene, rng)
        depth, rgb, inst = render(scene, T, H, W, f, device, seed=int(rng.integers(1 << 31)),
                                  use_kernel=torch.device(device).type == "cuda")
        masks, ids = instance_masks(inst, max_objects, min_pixels)
        if len(ids) == 0:
            continue
        frames.append({"rgb": rgb.cpu().numpy(), "depth": depth.cpu().numpy(), "masks": masks.cpu().numpy(),
                       "embs": query_embeddings(scene, ids, rng), "ids": ids, "T_wc": T})
    return frames
