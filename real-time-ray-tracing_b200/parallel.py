"""Multi-GPU host logic: one process and one RayTracer context per GPU, torch.distributed for the plumbing.

The path shards by IMAGE ROWS.  For the denoiser rank r owns the contiguous band rows [H*r/N, H*(r+1)/N) of every frame; the
path tracer splits the frame into interleaved STRIPS of `strip_height` rows, strip k going to rank k % N, because a frame is
not uniform (sky at the top, geometry at the bottom: contiguous bands leave half of the GPUs idle):

    build      every rank rebuilds the whole BVH redundantly (no communication), or rank 0 builds and the three traversal
               arrays (Triangle[], BVHNode[], TLAS BVHNode[]) are broadcast over NVLink — `bvh="rebuild" | "broadcast"`.
    pathtrace  each rank traces its strips (rtrt_set_row_strips), or its band when strip_height == 0 (rtrt_set_row_range).
    exchange   all-gather of the five G-buffers by rows (colour+mask 8, normal 8, albedo 8, depth 2, motion 4 = 30 B/px),
               so every rank holds the whole frame's denoiser inputs.  Strips are packed into one contiguous staging block
               per rank (row gather), all-gathered, and scattered back to their rows (one pack + one unpack kernel per buffer).
    denoise    each rank filters its band plus the 80-row dependency halo (csrc/denoise.cu, launch_denoise): the halo is
               recomputed instead of exchanged pass by pass, so there is no collective inside the denoiser.
    gather     all-gather of the band's denoised colour (the frame buffer), history colour and accumulation colour — what the
               next frame's temporal filters re-project from.

Progressive mode (BASELINE config 5): each rank accumulates N samples of its band in fp32 (rtrt_accumulate) and the bands are
gathered once at the end (16 B/px).

Everything here works on byte tensors and a process group, so the same code runs over NCCL on device buffers and over gloo
on CPU tensors (tests/test_parallel_cpu.py drives it with world_size 2 on the CPU).
"""
import numpy as np

from . import _lib as L


def row_bands(height, world):
    """Contiguous row bands, one per rank: [(begin, end), ...]; every row belongs to exactly one band."""
    if world < 1 or height < world:
        raise ValueError(f"cannot split {height} rows over {world} ranks")
    return [(height * r // world, height * (r + 1) // world) for r in range(world)]


def all_gather_rows(buf, row_bytes, bands, rank, group=None):
    """In-place all-gather of a row-major full-frame byte tensor `buf` (uint8, H*row_bytes): on entry rank r's rows
    bands[r] are valid, on return all rows are.  Equal bands -> one all_gather_into_tensor (in place: NCCL recognises the
    input as a slice of the output); ragged bands -> one broadcast per rank."""
    import torch.distributed as dist
    world = len(bands)
    if world == 1:
        return
    sizes = {e - b for b, e in bands}
    if len(sizes) == 1:
        b, e = bands[rank]
        mine = buf[b * row_bytes: e * row_bytes]
        dist.all_gather_into_tensor(buf[: bands[-1][1] * row_bytes], mine if buf.is_cuda else mine.clone(), group=group)
    else:
        for r, (b, e) in enumerate(bands):
            dist.broadcast(buf[b * row_bytes: e * row_bytes], src=dist.get_global_rank(group, r) if group is not None else r, group=group)


def strip_rows(height, strip_height, world):
    """Rows of each rank under the interleaved split: strip k = rows [k*h, (k+1)*h) belongs to rank k % world."""
    rows = np.arange(height)
    owner = (rows // strip_height) % world
    return [rows[owner == r] for r in range(world)]


class StripGather:
    """All-gather of interleaved row strips of row-major full-frame byte tensors.  Per buffer: index_select of the own rows into
    the rank's slot of a staging tensor, one all_gather_into_tensor, index_copy_ of all rows back.  Ranks with fewer rows pad
    their slot by repeating their last row (the duplicate is written twice with the same bytes)."""

    def __init__(self, height, strip_height, world, rank, device, group=None):
        import torch
        self.world, self.rank, self.group, self.H = world, rank, group, height
        per = strip_rows(height, strip_height, world)
        if any(len(p) == 0 for p in per):
            raise ValueError("strip split leaves a rank without rows")
        self.nmax = max(len(p) for p in per)
        padded = [np.concatenate([p, np.full(self.nmax - len(p), p[-1])]) for p in per]
        self.mine = torch.from_numpy(padded[rank].astype(np.int64)).to(device)
        self.all_rows = torch.from_numpy(np.concatenate(padded).astype(np.int64)).to(device)
        self._staging = {}

    def gather(self, buf, row_bytes):
        """buf: uint8 tensor of H*row_bytes; on entry the own strips are valid, on return every row is."""
        import torch
        import torch.distributed as dist
        if self.world == 1:
            return
        wide = torch.int64 if row_bytes % 8 == 0 and buf.data_ptr() % 8 == 0 else torch.uint8
        item = 8 if wide is torch.int64 else 1
        v = buf[: self.H * row_bytes].view(wide).view(self.H, row_bytes // item)
        key = (row_bytes, wide)
        if key not in self._staging:
            self._staging[key] = torch.empty((self.world * self.nmax, row_bytes // item), dtype=wide, device=buf.device)
        st = self._staging[key]
        mine = st[self.rank * self.nmax: (self.rank + 1) * self.nmax]
        torch.index_select(v, 0, self.mine, out=mine)
        dist.all_gather_into_tensor(st, mine if buf.is_cuda else mine.clone(), group=self.group)
        v.index_copy_(0, self.all_rows, st)


class _DevBytes:
    """A borrowed device buffer as a __cuda_array_interface__ object (so torch / NCCL can address it)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}


# (buffer id, bytes per pixel)
GBUFFERS = [(L.BUF_RENDER_COLOR, 8), (L.BUF_NORMAL, 8), (L.BUF_ALBEDO, 8), (L.BUF_DEPTH, 2), (L.BUF_MOTION_VECTOR, 4)]
DENOISED = [(L.BUF_RENDER_COLOR, 8), (L.BUF_HISTORY_COLOR, 8), (L.BUF_ACCUMULATION_COLOR, 8)]
BVH_ARRAYS = [L.BUF_TRIANGLES, L.BUF_BVH_NODES, L.BUF_TLAS_BVH_NODES]


class RowShardedRayTracer:
    """Drives one RayTracer context as rank `rank` of `world` row-sharded GPUs.  Collectives run on the context's stream."""

    def __init__(self, rt, width, height, rank, world, bvh="rebuild", group=None, strip_height=8):
        if bvh not in ("rebuild", "broadcast"):
            raise ValueError("bvh must be 'rebuild' or 'broadcast'")
        self.rt, self.W, self.H, self.rank, self.world, self.bvh, self.group = rt, width, height, rank, world, bvh, group
        self.bands = row_bands(height, world)
        rt.set_row_range(*self.bands[rank])
        self.strip_height = strip_height if world > 1 else 0
        self._strips = None
        if self.strip_height:
            rt.set_row_strips(self.strip_height, world, rank)
        self._stream = None

    @property
    def band(self):
        return self.bands[self.rank]

    def _on_stream(self):
        import torch
        if self._stream is None:
            self._stream = torch.cuda.ExternalStream(self.rt.stream)
        return torch.cuda.stream(self._stream)

    def _tensor(self, which):
        import torch
        ptr, nbytes = self.rt.get_buffer(which)
        return torch.as_tensor(_DevBytes(ptr, nbytes), device="cuda")

    def _gather(self, bufs, strips=False):
        """all-gather full-frame buffers by contiguous bands, or by the path tracer's interleaved strips"""
        if self.world == 1:
            return
        with self._on_stream():
            if strips and self.strip_height:
                if self._strips is None:
                    self._strips = StripGather(self.H, self.strip_height, self.world, self.rank, "cuda", self.group)
                for which, bpp in bufs:
                    self._strips.gather(self._tensor(which), self.W * bpp)
            else:
                for which, bpp in bufs:
                    all_gather_rows(self._tensor(which), self.W * bpp, self.bands, self.rank, self.group)

    # -- stages -------------------------------------------------------------------------------------------------------
    def build_bvh(self):
        """Rebuild on every rank, or build on rank 0 and broadcast the traversal arrays."""
        import torch.distributed as dist
        if self.bvh == "rebuild" or self.world == 1:
            self.rt.build_bvh()
            return
        if self.rank == 0:
            self.rt.build_bvh()
        with self._on_stream():
            for which in BVH_ARRAYS:
                dist.broadcast(self._tensor(which), src=0, group=self.group)
        if self.rank != 0:
            self.rt.set_option("assume_bvh_valid", 1)

    def pathtrace(self):
        self.rt.pathtrace()

    def exchange_gbuffers(self):
        self._gather(GBUFFERS, strips=True)

    def denoise(self):
        self.rt.denoise()

    def gather_frame(self):
        self._gather(DENOISED)

    def draw(self, frame_number=None):
        """One frame of RayTracer::draw, row-sharded; on return RTRT_BUF_RENDER_COLOR holds the whole denoised frame on
        every rank and the history buffers are complete for the next frame."""
        if frame_number is not None:
            self.rt.set_frame_number(frame_number)
        self.build_bvh()
        self.pathtrace()
        self.exchange_gbuffers()
        self.denoise()
        self.gather_frame()

    # -- progressive mode (BASELINE config 5) ---------------------------------------------------------------------------
    def render_progressive(self, spp, first_frame=1, partition="samples"):
        """spp path-traced samples accumulated in fp32 (rtrt_accumulate); on return RTRT_BUF_ACCUM_RADIANCE holds, on every rank, the
        sum over all spp samples in .xyz and the sample count in .w.
        partition="samples": rank r renders the WHOLE frame for the samples s with s % N == r, then one NCCL all-reduce (sum) of the
                             accumulation buffer (16 B/px).  Per-kernel latency tails do not shrink with the pixel count, so this
                             scales better than splitting every sample N ways (measured).
        partition="rows":    every rank renders its row strips of every sample, then one all-gather of the strips."""
        import torch
        import torch.distributed as dist
        if partition not in ("samples", "rows"):
            raise ValueError("partition must be 'samples' or 'rows'")
        self.build_bvh()
        if partition == "rows" or self.world == 1:
            for s in range(spp):
                self.rt.set_frame_number(first_frame + s)
                self.rt.pathtrace()
                self.rt.accumulate(reset=(s == 0))
            self._gather([(L.BUF_ACCUM_RADIANCE, 16)], strips=True)
            return
        self.rt.set_row_range(0, self.H); self.rt.set_row_strips(0, 1, 0)          # whole frames on every rank
        try:
            mine = [s for s in range(spp) if s % self.world == self.rank]
            for k, s in enumerate(mine):
                self.rt.set_frame_number(first_frame + s)
                self.rt.pathtrace()
                self.rt.accumulate(reset=(k == 0))
            if not mine:                                                            # more ranks than samples: contribute zeros
                self.rt.set_frame_number(first_frame); self.rt.pathtrace(); self.rt.accumulate(reset=True)
            with self._on_stream():
                t = self._tensor(L.BUF_ACCUM_RADIANCE).view(torch.float32)
                if not mine:
                    t.zero_()
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        finally:
            self.rt.set_row_range(*self.bands[self.rank])
            if self.strip_height:
                self.rt.set_row_strips(self.strip_height, self.world, self.rank)


def assemble_rows(parts, bands, row_items):
    """Host-side twin of the gather: stitch per-rank arrays (each full-frame sized, valid only on its band) into one."""
    out = np.array(parts[0], copy=True)
    for p, (b, e) in zip(parts, bands):
        out[b * row_items: e * row_items] = p[b * row_items: e * row_items]
    return out


# ------------------------------------------------------------------------------------------------------------------
# C-level multi-GPU (csrc/multi.cu): strips for the path tracer, bands + per-pass halo exchange for the denoiser, NCCL inside the library
# ------------------------------------------------------------------------------------------------------------------
def plan_rows(height, world, strip_height, rank, peer, kind, halo=0):
    """rtrt_multi_plan_rows (pure host function): the rows `rank` sends to `peer` in exchange kind 0 (G-buffer redistribution),
    1 (halo of `halo` rows), 2 (frame gather)."""
    import ctypes as C
    lib = L.load()
    out = np.zeros(height, np.int32)
    n = lib.rtrt_multi_plan_rows(height, world, strip_height, rank, peer, kind, halo, out.ctypes.data_as(C.c_void_p), height)
    if n < 0:
        raise ValueError(f"rtrt_multi_plan_rows: {L.STATUS_NAMES.get(n, n)}")
    return out[:n].copy()


def band_bounds(height, world):
    """The denoiser bands of csrc/multi.cu: whole 16-row tiles, [(begin, end), ...]."""
    tiles = (height + 15) // 16
    b = [min(height, (tiles * r // world) * 16) for r in range(world)] + [height]
    return list(zip(b[:-1], b[1:]))


class MultiGpuRayTracer:
    """Rank `rank` of `world` through the library's own multi-GPU entry points (rtrt_multi_init / rtrt_render_frame_multi).
    `unique_id`: the 128 bytes of rtrt_multi_unique_id made on rank 0 (use `MultiGpuRayTracer.bootstrap` under torch.distributed)."""

    def __init__(self, rt, rank, world, unique_id, mode="rows", strip_height=8):
        import ctypes as C
        self.rt, self.rank, self.world = rt, rank, world
        buf = (C.c_ubyte * 128).from_buffer_copy(bytes(unique_id))
        rt._check(rt._lib.rtrt_multi_init(rt._h, buf, rank, world))
        self.set_mode(mode, strip_height)

    @staticmethod
    def unique_id():
        import ctypes as C
        buf = (C.c_ubyte * 128)()
        rc = L.load().rtrt_multi_unique_id(buf)
        if rc:
            raise RuntimeError(f"rtrt_multi_unique_id: {L.STATUS_NAMES.get(rc, rc)} (NCCL not loadable?)")
        return bytes(buf)

    @classmethod
    def bootstrap(cls, rt, mode="rows", strip_height=8):
        """Under an initialised torch.distributed process group: rank 0 makes the NCCL id, broadcast_object_list hands it out."""
        import torch.distributed as dist
        rank, world = dist.get_rank(), dist.get_world_size()
        box = [cls.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        return cls(rt, rank, world, box[0], mode, strip_height)

    def set_mode(self, mode, strip_height=8):
        self.rt._check(self.rt._lib.rtrt_multi_set_mode(self.rt._h, {"rows": L.MULTI_ROWS, "samples": L.MULTI_SAMPLES}[mode], strip_height))
        self.mode = mode

    def update_vertices(self, pos_ptr=None, nrm_ptr=None):
        self.rt._check(self.rt._lib.rtrt_update_vertices_multi(self.rt._h, pos_ptr, nrm_ptr, self.rt._nverts))

    def draw(self, frame_number=None):
        if frame_number is not None:
            self.rt.set_frame_number(frame_number)
        self.rt._check(self.rt._lib.rtrt_render_frame_multi(self.rt._h))

    def render_progressive(self, spp, first_frame=1):
        self.rt._check(self.rt._lib.rtrt_render_progressive_multi(self.rt._h, spp, first_frame))

    def band(self):
        import ctypes as C
        b, e, x, n = C.c_int32(), C.c_int32(), C.c_uint64(), C.c_uint64()
        self.rt._check(self.rt._lib.rtrt_multi_get_band(self.rt._h, C.byref(b), C.byref(e), C.byref(x), C.byref(n)))
        return b.value, e.value, x.value, n.value


def link_local(rts):
    """rtrt_multi_link_local: the contexts of `rts` become ranks 0..N-1 of an in-process group (drive each from its own thread)."""
    import ctypes as C
    arr = (C.c_void_p * len(rts))(*[rt._h for rt in rts])
    rc = L.load().rtrt_multi_link_local(arr, len(rts))
    if rc:
        raise RuntimeError(f"rtrt_multi_link_local: {L.STATUS_NAMES.get(rc, rc)}")
