"""Synthetic procedural inputs for tests and benchmarks.  Pure NumPy, host side;
these only produce INPUT arrays (vertex / index / normal buffers, rays, camera blobs) — no hot-path
arithmetic lives here.

Terrain(gx, gz, seed): gx x gz quads (2 triangles each) over [-50,50]^2 in xz, height = 4-octave value
noise from a 32-bit integer hash (WangHash, src/blueNoiseRandGen.h:6-14 — no libm dependence), amplitude 8.
Triangle order = Morton order of the centroids (mirrors the offline pre-sort of tool/meshProcessor.cpp:146-200),
or a seeded shuffle to force an on-GPU global sort.
"""
import numpy as np


def wang_hash(a):
    a = np.asarray(a, dtype=np.uint32).copy()
    with np.errstate(over="ignore"):
        a = (a ^ np.uint32(61)) ^ (a >> np.uint32(16))
        a = a + (a << np.uint32(3))
        a = a ^ (a >> np.uint32(4))
        a = a * np.uint32(0x27D4EB2D)
        a = a ^ (a >> np.uint32(15))
    return a


def _lattice(ix, iz, seed):
    with np.errstate(over="ignore"):
        h = wang_hash(ix.astype(np.uint32) * np.uint32(73856093) ^ iz.astype(np.uint32) * np.uint32(19349663)
                      ^ np.uint32(seed) * np.uint32(83492791))
    return (h >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)   # [0,1), exact in fp32


def value_noise(x, z, seed):
    """Bilinear value noise with smoothstep fade; x,z float32 arrays in lattice units."""
    x0 = np.floor(x); z0 = np.floor(z)
    fx = (x - x0).astype(np.float32); fz = (z - z0).astype(np.float32)
    ix = x0.astype(np.int64); iz = z0.astype(np.int64)
    sx = fx * fx * (np.float32(3) - np.float32(2) * fx)
    sz = fz * fz * (np.float32(3) - np.float32(2) * fz)
    a = _lattice(ix, iz, seed); b = _lattice(ix + 1, iz, seed)
    c = _lattice(ix, iz + 1, seed); d = _lattice(ix + 1, iz + 1, seed)
    top = a + sx * (b - a)
    bot = c + sx * (d - c)
    return (top + sz * (bot - top)).astype(np.float32)


def expand10(v):
    v = v.astype(np.uint32)
    v = (v | (v << np.uint32(16))) & np.uint32(0x030000FF)
    v = (v | (v << np.uint32(8))) & np.uint32(0x0300F00F)
    v = (v | (v << np.uint32(4))) & np.uint32(0x030C30C3)
    v = (v | (v << np.uint32(2))) & np.uint32(0x09249249)
    return v


def morton60(cx, cy, cz, lo, hi):
    """60-bit code of tool/meshProcessor.cpp:35-62 (two 30-bit codes of the low / high 10 bits per axis)."""
    scale = np.float32((1 << 20) - 1)
    def q(c, l, h):
        ext = np.float32(h) - np.float32(l)
        u = (c - np.float32(l)) / (ext if ext > 0 else np.float32(1))
        return np.clip(u * scale, 0, scale).astype(np.uint32)
    x, y, z = q(cx, lo[0], hi[0]), q(cy, lo[1], hi[1]), q(cz, lo[2], hi[2])
    def m30(a, b, c):
        return (expand10(a) | (expand10(b) << np.uint32(1)) | (expand10(c) << np.uint32(2))).astype(np.uint64)
    lo30 = m30(x & 1023, y & 1023, z & 1023)
    hi30 = m30(x >> 10, y >> 10, z >> 10)
    return (hi30 << np.uint64(30)) | lo30


def morton63(cx, cy, cz, lo, hi):
    """63-bit code of the product's "morton_bits" = 63 mode: 21 bits per axis (same float32 quantisation rule as morton60 with
    2^21 - 1), bit i of x / y / z at position 3 i / 3 i + 1 / 3 i + 2."""
    scale = np.float32((1 << 21) - 1)
    def q(c, l, h):
        ext = np.float32(h) - np.float32(l)
        u = (c - np.float32(l)) / (ext if ext > 0 else np.float32(1))
        return np.clip(u * scale, 0, scale).astype(np.uint64)
    code = np.zeros(np.asarray(cx).shape, np.uint64)
    for a, v in enumerate((q(cx, lo[0], hi[0]), q(cy, lo[1], hi[1]), q(cz, lo[2], hi[2]))):
        for i in range(21):
            code |= ((v >> np.uint64(i)) & np.uint64(1)) << np.uint64(3 * i + a)
    return code


def terrain(gx, gz, seed=1, order="morton", extent=50.0, amplitude=8.0, shuffle_seed=1234):
    """Returns (pos [V,3] f32, nrm [V,3] f32, idx [N,3] u32) with N = 2*gx*gz."""
    xs = np.linspace(-extent, extent, gx + 1, dtype=np.float32)
    zs = np.linspace(-extent, extent, gz + 1, dtype=np.float32)
    X, Z = np.meshgrid(xs, zs, indexing="ij")
    h = np.zeros_like(X, dtype=np.float32)
    freq, amp = np.float32(0.04), np.float32(amplitude)
    for o in range(4):
        h += amp * (value_noise(X * freq + np.float32(100.0), Z * freq + np.float32(100.0), seed + o) - np.float32(0.5))
        freq *= np.float32(2.0); amp *= np.float32(0.5)
    pos = np.stack([X, h, Z], axis=-1).reshape(-1, 3).astype(np.float32)
    # normals from central differences of the height field (inputs only; any unit-ish normal works)
    dx = np.float32(2 * extent / gx); dz = np.float32(2 * extent / gz)
    dhdx = np.gradient(h, axis=0) / dx
    dhdz = np.gradient(h, axis=1) / dz
    n = np.stack([-dhdx, np.ones_like(h), -dhdz], axis=-1).astype(np.float32)
    n /= np.linalg.norm(n, axis=-1, keepdims=True).astype(np.float32)
    nrm = n.reshape(-1, 3).astype(np.float32)
    ii, jj = np.meshgrid(np.arange(gx, dtype=np.int64), np.arange(gz, dtype=np.int64), indexing="ij")
    v00 = ii * (gz + 1) + jj; v10 = v00 + (gz + 1); v01 = v00 + 1; v11 = v10 + 1
    t0 = np.stack([v00, v10, v11], axis=-1); t1 = np.stack([v00, v11, v01], axis=-1)
    idx = np.stack([t0, t1], axis=2).reshape(-1, 3).astype(np.uint32)
    if order == "morton":
        c = (pos[idx[:, 0]] + pos[idx[:, 1]] + pos[idx[:, 2]]) / np.float32(3.0)
        lo = pos.min(axis=0); hi = pos.max(axis=0)
        code = morton60(c[:, 0], c[:, 1], c[:, 2], lo, hi)
        idx = idx[np.argsort(code, kind="stable")]
    elif order == "shuffle":
        rng = np.random.default_rng(shuffle_seed)
        idx = idx[rng.permutation(idx.shape[0])]
    elif order != "grid":
        raise ValueError(order)
    return pos, nrm, np.ascontiguousarray(idx)


# named sizes of SURVEY.md §8 (all multiples of 1024 triangles)
TERRAIN_SIZES = {
    "C1": (64, 80),        # 10 240 tris
    "C2ref": (512, 1023),  # 1 047 552 tris = 1023 batches (largest the reference accepts)
    "C2": (512, 1024),     # 1 048 576 tris
    "C3": (1024, 2048),    # 4 194 304 tris
    "C5": (2048, 4096),    # 16 777 216 tris
}


def make_camera(pos, yaw, pitch, width, height, focal=5.0, aperture=1e-3, fov_x_deg=90.0):
    """Camera blob (src/kernel.cuh:78-101, 224 bytes) with the derived fields of Camera::update()
    (src/kernel.cuh:103-121) computed here on the host, exactly like the reference does (host libm)."""
    f = np.float32
    cam = np.zeros(56, dtype=np.float32)
    yaw, pitch = f(yaw), f(pitch)
    d = np.array([np.sin(yaw) * np.cos(pitch), np.sin(pitch), np.cos(yaw) * np.cos(pitch)], dtype=np.float32)
    res = np.array([width, height], dtype=np.float32)
    inv = (f(1.0) / res).astype(np.float32)
    fovx = f(fov_x_deg * np.pi / 180.0)
    fovy = f(fovx / res[0] * res[1])
    thf = np.array([np.tan(fovx / f(2)), np.tan(fovy / f(2))], dtype=np.float32)
    up0 = np.array([0, 1, 0], dtype=np.float32)
    left = np.cross(up0, d).astype(np.float32); left /= f(np.linalg.norm(left))
    up = np.cross(d, left).astype(np.float32); up /= f(np.linalg.norm(up))
    cam[0:3] = np.asarray(pos, dtype=np.float32); cam[3] = pitch
    cam[4:7] = d; cam[7] = f(focal)
    cam[8:11] = left; cam[11] = f(aperture)
    cam[12:15] = up; cam[15] = yaw
    cam[16:18] = res; cam[18:20] = inv
    cam[20:22] = [fovx, fovy]; cam[22:24] = thf
    cam[24:27] = left * thf[0] * f(focal)          # adjustedLeft
    cam[28:31] = up * thf[1] * f(focal)            # adjustedUp
    cam[32:35] = d * f(focal)                      # adjustedFront
    cam[36:39] = left * f(aperture)                # apertureLeft
    cam[40:43] = up * f(aperture)                  # apertureUp
    return cam[:44].copy()


def primary_rays(cam, width, height):
    """Pinhole rays through pixel centres (GenerateRay, src/raygen.cuh:7-38, with offset 0.5 and no lens sample)."""
    from .api import RAY_DTYPE
    f = np.float32
    px, py = np.meshgrid(np.arange(width, dtype=np.float32), np.arange(height, dtype=np.float32), indexing="xy")
    u = ((px + f(0.5)) * cam[18]) * f(-2.0) + f(1.0)
    v = ((py + f(0.5)) * cam[19]) * f(-2.0) + f(1.0)
    p = cam[32:35][None, None, :] + cam[24:27][None, None, :] * u[..., None] + cam[28:31][None, None, :] * v[..., None]
    p = p.astype(np.float32)
    d = p / np.sqrt((p * p).sum(-1, keepdims=True)).astype(np.float32)
    rays = np.zeros(width * height, dtype=RAY_DTYPE)
    rays["orig"] = cam[0:3]
    rays["dir"] = d.reshape(-1, 3).astype(np.float32)
    return rays
