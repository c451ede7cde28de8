// Multi-GPU frame at the C level (SURVEY 8(b) rtrt_create_multi, 8(e)): one rtrt_ctx per GPU and rank, NCCL over NVLink for the three
// exchanges north_star names — (BVH inputs) vertex broadcast, the a-trous HALO EXCHANGE and the framebuffer gather — plus the
// redistribution of the path tracer's row strips to the denoiser's row bands.  No counterpart in the reference (it is single-GPU); the
// stages it drives are RayTracer::draw's (src/kernel.cu:259-362).
//
//   partition   path trace: interleaved strips of `strip_height` rows, strip k -> rank k % N (a frame is sky at the top and geometry at
//               the bottom: contiguous bands leave half of the GPUs idle).  Denoise: contiguous bands of whole 16-row tiles, so every
//               16x16 noise-gate tile (src/temporalDenoising.cuh:33-102) belongs to exactly one rank.
//   G           after the path trace every rank needs the five G-buffers (30 B/px) on its band +- 32 rows: each rank packs the rows it
//               traced that a peer needs into one block per peer (ONE pack kernel), grouped ncclSend / ncclRecv, one unpack kernel.
//   halo        before every denoise pass that reads neighbouring rows (7x7: 3 rows; dilated 5x5: 2 S = 6 / 12 / 24 rows, or 2 / 4 / 8 / 16 / 32
//               with the 5-iteration schedule; TemporalFilter2: 1 row) the ranks exchange that many rows of the CURRENT colour image with
//               the neighbouring bands (src/temporalDenoising.cuh:317-608 are the filters whose footprints these are).  Nothing is recomputed.
//   F           the band's denoised colour and accumulation colour (16 B/px) go to every rank: the frame buffer gather; it also completes the
//               temporal history, which the next frame's filters re-project from at motion-vector distance (any row).
//   samples     alternative partition (north_star: "by tile or by sample"): every rank renders the whole frame with its own sample index and the
//               denoised frames are averaged with one ncclAllReduce (rgb lanes only; the material mask lane is kept).
// Transports: NCCL (loaded with dlopen at the first use, so that a process that already holds an NCCL — e.g. through torch — shares it) and
// an in-process transport for N contexts driven by N host threads (rtrt_multi_link_local: copies between the contexts' staging buffers,
// host barriers) — the same exchange plans and kernels, used by the single-GPU tests and by single-process multi-GPU hosts.
#include "context.h"
#include <nccl.h>
#include <dlfcn.h>
#include <cuda_fp16.h>
#include <condition_variable>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

namespace rtrt {

int launch_denoise(Context& c);

// ------------------------------------------------------------------------------------------------------------------ NCCL through dlopen
struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;
    bool ok = false;
};

static NcclApi& nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);              // the copy this process already uses (torch's), if any
        if (!h) { const char* e = getenv("RTRT_NCCL_LIB"); if (e && *e) h = dlopen(e, RTLD_NOW | RTLD_GLOBAL); }
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) { api.error = std::string("cannot load libnccl.so.2 (set RTRT_NCCL_LIB): ") + (dlerror() ? dlerror() : ""); return; }
        bool all = true;
        auto sym = [&](const char* n) { void* p = dlsym(h, n); if (!p) { all = false; api.error = std::string("libnccl lacks ") + n; } return p; };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
        api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
        api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
        api.Broadcast = reinterpret_cast<decltype(api.Broadcast)>(sym("ncclBroadcast"));
        api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
        api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        api.ok = all;
    });
    return api;
}

#define RTRT_NCCL(ctx, call)                                                                                  \
    do {                                                                                                      \
        ncclResult_t _r = (call);                                                                             \
        if (_r != ncclSuccess) {                                                                              \
            char _b[384];                                                                                     \
            snprintf(_b, sizeof(_b), "NCCL error %d (%s) in %s", (int)_r, nccl_api().GetErrorString(_r), #call); \
            return (ctx).fail(RTRT_ERR_NCCL, _b);                                                             \
        }                                                                                                     \
    } while (0)

// ------------------------------------------------------------------------------------------------------------------ partition + plans (host)
enum { X_GBUF = 0, X_HALO = 1, X_FRAME = 2 };
#ifndef RTRT_MULTI_DIRECT
#define RTRT_MULTI_DIRECT 1     // 0: every exchange through the pack / unpack kernels (A/B)
#endif
constexpr int kGbufReach = 32;                   // rows around the band on which the denoiser reads G-buffers (guide reach of the widest dilated pass)

// bands of whole 16-row tiles; begin[r] .. begin[r + 1]
static void band_bounds(int height, int world, std::vector<int>& begin) {
    const int tiles = (height + 15) / 16;
    begin.resize(world + 1);
    for (int r = 0; r <= world; ++r) { const int b = (int)((long long)tiles * r / world) * 16; begin[r] = b < height ? b : height; }
    begin[world] = height;
}

// rows `rank` sends to `peer` in exchange (kind, halo): rows `rank` owns that `peer` needs and does not own itself
static void plan_rows(int height, int world, int stripH, int rank, int peer, int kind, int halo, std::vector<int>& rows) {
    rows.clear();
    if (rank == peer) return;
    std::vector<int> bb;
    band_bounds(height, world, bb);
    for (int y = 0; y < height; ++y) {
        bool own, need;
        if (kind == X_GBUF) {
            const int owner = (y / stripH) % world;
            own = owner == rank;
            need = y >= bb[peer] - kGbufReach && y < bb[peer + 1] + kGbufReach && owner != peer;
        } else if (kind == X_HALO) {
            own = y >= bb[rank] && y < bb[rank + 1];
            need = (y >= bb[peer] - halo && y < bb[peer]) || (y >= bb[peer + 1] && y < bb[peer + 1] + halo);
        } else {
            own = y >= bb[rank] && y < bb[rank + 1];
            need = !(y >= bb[peer] && y < bb[peer + 1]);
        }
        if (own && need) rows.push_back(y);
    }
}

struct XBufs { unsigned char* ptr[5]; int rowBytes[5]; int n; int rowBytesAll; };

struct XPlan {
    std::vector<int> sendStart, recvStart;       // world + 1 offsets into the concatenated row lists
    std::vector<int> sendFirst, recvFirst;       // first row of every peer's block
    bool contiguous = false;                     // every block is ONE range of consecutive rows (halo and frame exchanges: bands are contiguous):
                                                 // the rows then travel straight between the images, without the pack / unpack kernels
    int* dSendRows = nullptr; int* dRecvRows = nullptr;
    int sendTotal = 0, recvTotal = 0;
};

struct LocalGroup;

struct Multi {
    int rank = 0, world = 1;
    int mode = RTRT_MULTI_ROWS, stripH = 8;
    ncclComm_t comm = nullptr; bool ownsComm = false;
    LocalGroup* local = nullptr;
    std::vector<int> bandBegin;
    std::map<int, XPlan> plans;
    unsigned char* sendStage = nullptr; unsigned char* recvStage = nullptr; size_t stageBytes = 0;
    XPlan* current = nullptr; size_t curRowBytes = 0;   // plan / packed row size of the exchange in flight (read by the peers of the in-process transport)
    XBufs curBufs{};                             // ... and its images (direct exchanges: the peers copy out of them)
    int width = 0, height = 0;                   // frame size the plans were made for
    unsigned long long exchanges = 0, bytesSent = 0;
};

struct LocalGroup {
    int n = 0;
    std::vector<Context*> members;
    std::mutex mu; std::condition_variable cv; int waiting = 0; unsigned long long generation = 0;
    int refs = 0;
    void barrier() {
        std::unique_lock<std::mutex> lk(mu);
        const unsigned long long g = generation;
        if (++waiting == n) { waiting = 0; ++generation; cv.notify_all(); }
        else cv.wait(lk, [&] { return generation != g; });
    }
};

static Multi* multi_of(Context& c) { return reinterpret_cast<Multi*>(c.multi); }

static void free_plans(Multi& m) {
    for (auto& kv : m.plans) { if (kv.second.dSendRows) cudaFree(kv.second.dSendRows); if (kv.second.dRecvRows) cudaFree(kv.second.dRecvRows); }
    m.plans.clear();
}

static int get_plan(Context& c, Multi& m, int kind, int halo, XPlan** out) {
    if (m.width != c.width || m.height != c.height) { free_plans(m); m.width = c.width; m.height = c.height; band_bounds(c.height, m.world, m.bandBegin); }
    const int key = kind * 4096 + halo;
    auto it = m.plans.find(key);
    if (it == m.plans.end()) {
        XPlan p;
        std::vector<int> send, recv, rows;
        p.sendStart.assign(m.world + 1, 0); p.recvStart.assign(m.world + 1, 0);
        p.sendFirst.assign(m.world, 0); p.recvFirst.assign(m.world, 0);
        p.contiguous = kind != X_GBUF;
        auto one_range = [](const std::vector<int>& r) { return r.empty() || r.back() - r.front() + 1 == (int)r.size(); };
        for (int d = 0; d < m.world; ++d) {
            plan_rows(c.height, m.world, m.stripH, m.rank, d, kind, halo, rows);
            send.insert(send.end(), rows.begin(), rows.end()); p.sendStart[d + 1] = (int)send.size();
            p.sendFirst[d] = rows.empty() ? 0 : rows.front(); p.contiguous = p.contiguous && one_range(rows);
            plan_rows(c.height, m.world, m.stripH, d, m.rank, kind, halo, rows);
            recv.insert(recv.end(), rows.begin(), rows.end()); p.recvStart[d + 1] = (int)recv.size();
            p.recvFirst[d] = rows.empty() ? 0 : rows.front(); p.contiguous = p.contiguous && one_range(rows);
        }
        p.sendTotal = (int)send.size(); p.recvTotal = (int)recv.size();
        if (p.sendTotal) {
            RTRT_CUDA(c, cudaMalloc((void**)&p.dSendRows, send.size() * sizeof(int)));
            RTRT_CUDA(c, cudaMemcpyAsync(p.dSendRows, send.data(), send.size() * sizeof(int), cudaMemcpyHostToDevice, c.stream));
        }
        if (p.recvTotal) {
            RTRT_CUDA(c, cudaMalloc((void**)&p.dRecvRows, recv.size() * sizeof(int)));
            RTRT_CUDA(c, cudaMemcpyAsync(p.dRecvRows, recv.data(), recv.size() * sizeof(int), cudaMemcpyHostToDevice, c.stream));
        }
        RTRT_CUDA(c, cudaStreamSynchronize(c.stream));       // the host vectors go out of scope
        it = m.plans.emplace(key, std::move(p)).first;
    }
    *out = &it->second;
    return RTRT_OK;
}

// ------------------------------------------------------------------------------------------------------------------ pack / unpack kernels
// One CTA per packed row: row k of the list is copied between the frame buffers (row-major, full width) and the staging block, where the
// rows of all buffers of one frame row lie back to back (rowBytesAll per row).  16-byte words (all row sizes are multiples of 16 when W % 8 == 0).
template <bool PACK, typename WORD>
__global__ void __launch_bounds__(256) k_x_rows(const XBufs b, const int* __restrict__ rows, unsigned char* stage) {
    const int k = blockIdx.x, y = rows[k];
    unsigned char* st = stage + (size_t)k * b.rowBytesAll;
    int off = 0;
    for (int j = 0; j < b.n; ++j) {
        WORD* img = reinterpret_cast<WORD*>(b.ptr[j] + (size_t)y * b.rowBytes[j]);
        WORD* s = reinterpret_cast<WORD*>(st + off);
        const int words = b.rowBytes[j] / (int)sizeof(WORD);
        for (int i = threadIdx.x; i < words; i += 256) { if (PACK) s[i] = img[i]; else img[i] = s[i]; }
        off += b.rowBytes[j];
    }
}

static int ensure_stage(Context& c, Multi& m, size_t bytes) {
    if (bytes <= m.stageBytes) return RTRT_OK;
    if (m.local) { RTRT_CUDA(c, cudaStreamSynchronize(c.stream)); }
    if (m.sendStage) cudaFree(m.sendStage);
    if (m.recvStage) cudaFree(m.recvStage);
    m.sendStage = m.recvStage = nullptr; m.stageBytes = 0;
    RTRT_CUDA(c, cudaMalloc((void**)&m.sendStage, bytes)); RTRT_CUDA(c, cudaMalloc((void**)&m.recvStage, bytes));
    m.stageBytes = bytes;
    return RTRT_OK;
}

// peers' blocks: NCCL = grouped send / recv on the context stream; in-process = barrier, copies out of the peers' send staging, barrier
static int transport(Context& c, Multi& m, const std::vector<size_t>& sendOff, const std::vector<size_t>& recvOff) {
    if (m.local) {
        RTRT_CUDA(c, cudaStreamSynchronize(c.stream));                      // my blocks are packed
        m.local->barrier();                                                 // everybody's are
        for (int s = 0; s < m.world; ++s) {
            const size_t bytes = recvOff[s + 1] - recvOff[s];
            if (s == m.rank || bytes == 0) continue;
            Context* pc = m.local->members[s];
            Multi* pm = multi_of(*pc);
            // the peer's block for me starts at ITS send offset of rank `me`; both sides derive the sizes from the same plan formula
            const size_t peerOff = (size_t)pm->current->sendStart[m.rank] * m.curRowBytes;
            RTRT_CUDA(c, cudaMemcpyPeerAsync(m.recvStage + recvOff[s], c.device, pm->sendStage + peerOff, pc->device, bytes, c.stream));
        }
        RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
        m.local->barrier();                                                 // send staging may be overwritten again
        return RTRT_OK;
    }
    NcclApi& N = nccl_api();
    RTRT_NCCL(c, N.GroupStart());
    for (int d = 0; d < m.world; ++d) {
        if (d == m.rank) continue;
        const size_t sb = sendOff[d + 1] - sendOff[d], rb = recvOff[d + 1] - recvOff[d];
        if (sb) RTRT_NCCL(c, N.Send(m.sendStage + sendOff[d], sb, ncclUint8, d, m.comm, c.stream));
        if (rb) RTRT_NCCL(c, N.Recv(m.recvStage + recvOff[d], rb, ncclUint8, d, m.comm, c.stream));
    }
    RTRT_NCCL(c, N.GroupEnd());
    return RTRT_OK;
}

// contiguous blocks: rows go straight from the sender's image into the same rows of the receiver's image
static int exchange_direct(Context& c, Multi& m, XPlan& p, const XBufs& bufs) {
    m.current = &p; m.curBufs = bufs;
    size_t sent = 0;
    if (m.local) {
        RTRT_CUDA(c, cudaStreamSynchronize(c.stream));                      // my rows are computed
        m.local->barrier();
        for (int s = 0; s < m.world; ++s) {
            const int rows = p.recvStart[s + 1] - p.recvStart[s];
            if (s == m.rank || rows == 0) continue;
            Context* pc = m.local->members[s];
            const XBufs& pb = multi_of(*pc)->curBufs;
            for (int j = 0; j < bufs.n; ++j) {
                const size_t off = (size_t)p.recvFirst[s] * bufs.rowBytes[j], bytes = (size_t)rows * bufs.rowBytes[j];
                RTRT_CUDA(c, cudaMemcpyPeerAsync(bufs.ptr[j] + off, c.device, pb.ptr[j] + off, pc->device, bytes, c.stream));
            }
        }
        RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
        m.local->barrier();
    } else {
        NcclApi& N = nccl_api();
        RTRT_NCCL(c, N.GroupStart());
        for (int d = 0; d < m.world; ++d) {
            if (d == m.rank) continue;
            const int sr = p.sendStart[d + 1] - p.sendStart[d], rr = p.recvStart[d + 1] - p.recvStart[d];
            for (int j = 0; j < bufs.n; ++j) {
                if (sr) RTRT_NCCL(c, N.Send(bufs.ptr[j] + (size_t)p.sendFirst[d] * bufs.rowBytes[j], (size_t)sr * bufs.rowBytes[j], ncclUint8, d, m.comm, c.stream));
                if (rr) RTRT_NCCL(c, N.Recv(bufs.ptr[j] + (size_t)p.recvFirst[d] * bufs.rowBytes[j], (size_t)rr * bufs.rowBytes[j], ncclUint8, d, m.comm, c.stream));
            }
        }
        RTRT_NCCL(c, N.GroupEnd());
    }
    for (int j = 0; j < bufs.n; ++j) sent += (size_t)p.sendTotal * bufs.rowBytes[j];
    m.exchanges++; m.bytesSent += sent;
    return RTRT_OK;
}

static int exchange(Context& c, Multi& m, int kind, int halo, XBufs bufs) {
    XPlan* p = nullptr;
    { int rc = get_plan(c, m, kind, halo, &p); if (rc) return rc; }
    if (p->contiguous && RTRT_MULTI_DIRECT) return exchange_direct(c, m, *p, bufs);
    bufs.rowBytesAll = 0;
    bool wide16 = true;
    for (int j = 0; j < bufs.n; ++j) { bufs.rowBytesAll += bufs.rowBytes[j]; wide16 = wide16 && bufs.rowBytes[j] % 16 == 0 && (reinterpret_cast<uintptr_t>(bufs.ptr[j]) & 15) == 0; }
    const size_t need = (size_t)(p->sendTotal > p->recvTotal ? p->sendTotal : p->recvTotal) * bufs.rowBytesAll;
    { int rc = ensure_stage(c, m, need ? need : 16); if (rc) return rc; }
    m.current = p; m.curRowBytes = (size_t)bufs.rowBytesAll;
    if (p->sendTotal) {
        if (wide16) k_x_rows<true, uint4><<<p->sendTotal, 256, 0, c.stream>>>(bufs, p->dSendRows, m.sendStage);
        else        k_x_rows<true, uint16_t><<<p->sendTotal, 256, 0, c.stream>>>(bufs, p->dSendRows, m.sendStage);
        c.launches++;
    }
    std::vector<size_t> so(m.world + 1), ro(m.world + 1);
    for (int d = 0; d <= m.world; ++d) { so[d] = (size_t)p->sendStart[d] * bufs.rowBytesAll; ro[d] = (size_t)p->recvStart[d] * bufs.rowBytesAll; }
    { int rc = transport(c, m, so, ro); if (rc) return rc; }
    if (p->recvTotal) {
        if (wide16) k_x_rows<false, uint4><<<p->recvTotal, 256, 0, c.stream>>>(bufs, p->dRecvRows, m.recvStage);
        else        k_x_rows<false, uint16_t><<<p->recvTotal, 256, 0, c.stream>>>(bufs, p->dRecvRows, m.recvStage);
        c.launches++;
    }
    m.exchanges++; m.bytesSent += so[m.world];
    RTRT_CUDA(c, cudaGetLastError());
    return RTRT_OK;
}

// the denoiser's hook: `rows` halo rows of the colour image the next pass reads
static int halo_hook(Context& c, void* colour, int rows) {
    Multi* m = multi_of(c);
    if (!m || m->world == 1) return RTRT_OK;
    XBufs b{}; b.n = 1; b.ptr[0] = reinterpret_cast<unsigned char*>(colour); b.rowBytes[0] = c.width * 8;
    return exchange(c, *m, X_HALO, rows, b);
}

// ------------------------------------------------------------------------------------------------------------------ sample-partition combine
// rgb lanes of the colour image (half3 + u16 mask) <-> a half4 staging image: the all-reduce must not average the mask lane's bit patterns
__global__ void __launch_bounds__(256) k_rgb_to_stage(const ushort4* __restrict__ colour, ushort4* __restrict__ stage, size_t n) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i < n) { ushort4 p = colour[i]; p.w = 0; stage[i] = p; }
}
__global__ void __launch_bounds__(256) k_stage_to_rgb(ushort4* __restrict__ colour, const ushort4* __restrict__ stage, size_t n) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i < n) { const ushort4 s = stage[i]; ushort4 p = colour[i]; p.x = s.x; p.y = s.y; p.z = s.z; colour[i] = p; }
}
// in-process transport: mean (half) / sum (float) over the N staging images gathered into recvStage
__global__ void __launch_bounds__(256) k_mean_half(const __half* __restrict__ all, __half* __restrict__ out, size_t n, int world) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    float s = 0.0f;
    for (int r = 0; r < world; ++r) s += __half2float(all[(size_t)r * n + i]);
    out[i] = __float2half_rn(s / (float)world);
}
__global__ void __launch_bounds__(256) k_sum_float(const float* __restrict__ all, float* __restrict__ out, size_t n, int world) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    float s = 0.0f;
    for (int r = 0; r < world; ++r) s += all[(size_t)r * n + i];
    out[i] = s;
}

// all-reduce of `bytes` at `buf` (in place): mean of halves, or sum of floats
static int all_reduce(Context& c, Multi& m, void* buf, size_t bytes, bool halfMean) {
    if (m.world == 1) return RTRT_OK;
    if (!m.local) {
        NcclApi& N = nccl_api();
        if (halfMean) RTRT_NCCL(c, N.AllReduce(buf, buf, bytes / 2, ncclHalf, ncclAvg, m.comm, c.stream));
        else          RTRT_NCCL(c, N.AllReduce(buf, buf, bytes / 4, ncclFloat, ncclSum, m.comm, c.stream));
        return RTRT_OK;
    }
    { int rc = ensure_stage(c, m, bytes * (size_t)m.world); if (rc) return rc; }
    RTRT_CUDA(c, cudaMemcpyAsync(m.sendStage, buf, bytes, cudaMemcpyDeviceToDevice, c.stream));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    m.local->barrier();
    for (int s = 0; s < m.world; ++s) {
        Context* pc = m.local->members[s];
        RTRT_CUDA(c, cudaMemcpyPeerAsync(m.recvStage + (size_t)s * bytes, c.device, multi_of(*pc)->sendStage, pc->device, bytes, c.stream));
    }
    if (halfMean) k_mean_half<<<(unsigned)((bytes / 2 + 255) / 256), 256, 0, c.stream>>>(reinterpret_cast<const __half*>(m.recvStage), reinterpret_cast<__half*>(buf), bytes / 2, m.world);
    else          k_sum_float<<<(unsigned)((bytes / 4 + 255) / 256), 256, 0, c.stream>>>(reinterpret_cast<const float*>(m.recvStage), reinterpret_cast<float*>(buf), bytes / 4, m.world);
    c.launches++;
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    m.local->barrier();
    return RTRT_OK;
}

static int broadcast0(Context& c, Multi& m, void* buf, size_t bytes) {
    if (m.world == 1) return RTRT_OK;
    if (!m.local) { RTRT_NCCL(c, nccl_api().Broadcast(buf, buf, bytes, ncclUint8, 0, m.comm, c.stream)); return RTRT_OK; }
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    m.local->barrier();
    if (m.rank != 0) {
        Context* p0 = m.local->members[0];
        const unsigned char* src = buf == c.dPos ? reinterpret_cast<unsigned char*>(p0->dPos) : reinterpret_cast<unsigned char*>(p0->dNrm);
        RTRT_CUDA(c, cudaMemcpyPeerAsync(buf, c.device, src, p0->device, bytes, c.stream));
        RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    }
    m.local->barrier();
    return RTRT_OK;
}

static void apply_partition(Context& c, Multi& m) {
    band_bounds(c.height, m.world, m.bandBegin);
    if (m.mode == RTRT_MULTI_ROWS && m.world > 1) {
        c.rowBegin = m.bandBegin[m.rank]; c.rowEnd = m.bandBegin[m.rank + 1];
        c.stripH = m.stripH; c.stripShards = m.world; c.stripShard = m.rank;
        c.bandHaloRows = 0; c.haloFn = halo_hook;
    } else {
        c.rowBegin = 0; c.rowEnd = c.height; c.stripH = 0; c.stripShards = 1; c.stripShard = 0;
        c.bandHaloRows = -1; c.haloFn = nullptr;
    }
}

static int attach(Context& c, Multi* m) {
    if (c.multi) return c.fail(RTRT_ERR_STATE, "rtrt_multi: the context already belongs to a multi-GPU group");
    if ((c.height + 15) / 16 < m->world) return c.fail(RTRT_ERR_CAPACITY, "rtrt_multi: fewer 16-row tiles than ranks");
    if (c.width % 2) return c.fail(RTRT_ERR_ARG, "rtrt_multi: the frame width must be even");
    c.multi = m;
    apply_partition(c, *m);
    return RTRT_OK;
}

void multi_release(Context& c) {              // rtrt_destroy
    Multi* m = multi_of(c);
    if (!m) return;
    free_plans(*m);
    if (m->sendStage) cudaFree(m->sendStage);
    if (m->recvStage) cudaFree(m->recvStage);
    if (m->comm && m->ownsComm && nccl_api().ok) nccl_api().CommDestroy(m->comm);
    if (m->local) { std::unique_lock<std::mutex> lk(m->local->mu); const bool last = --m->local->refs == 0; lk.unlock(); if (last) delete m->local; }
    delete m;
    c.multi = nullptr; c.haloFn = nullptr; c.bandHaloRows = -1;
}

}  // namespace rtrt

using namespace rtrt;

extern "C" {

int rtrt_multi_unique_id(void* id_out) {
    if (!id_out) return RTRT_ERR_ARG;
    NcclApi& N = nccl_api();
    if (!N.ok) return RTRT_ERR_NCCL;
    static_assert(sizeof(ncclUniqueId) == RTRT_NCCL_UNIQUE_ID_BYTES, "ncclUniqueId size");
    ncclUniqueId id;
    if (N.GetUniqueId(&id) != ncclSuccess) return RTRT_ERR_NCCL;
    std::memcpy(id_out, &id, sizeof(id));
    return RTRT_OK;
}

int rtrt_multi_init(rtrt_ctx* ctx, const void* unique_id, int rank, int world) {
    if (!ctx) return RTRT_ERR_ARG;
    Context& c = ctx->c;
    if (!unique_id || world < 1 || rank < 0 || rank >= world) return c.fail(RTRT_ERR_ARG, "rtrt_multi_init: bad rank / world / id");
    NcclApi& N = nccl_api();
    if (!N.ok) return c.fail(RTRT_ERR_NCCL, N.error.c_str());
    RTRT_CUDA(c, cudaSetDevice(c.device));
    Multi* m = new Multi(); m->rank = rank; m->world = world;
    ncclUniqueId id; std::memcpy(&id, unique_id, sizeof(id));
    ncclResult_t r = N.CommInitRank(&m->comm, world, id, rank);
    if (r != ncclSuccess) { delete m; char b[256]; snprintf(b, sizeof(b), "ncclCommInitRank failed: %s", N.GetErrorString(r)); return c.fail(RTRT_ERR_NCCL, b); }
    m->ownsComm = true;
    int rc = attach(c, m);
    if (rc) { N.CommDestroy(m->comm); delete m; }
    return rc;
}

int rtrt_multi_attach_comm(rtrt_ctx* ctx, void* nccl_comm, int rank, int world) {
    if (!ctx) return RTRT_ERR_ARG;
    Context& c = ctx->c;
    if (!nccl_comm || world < 1 || rank < 0 || rank >= world) return c.fail(RTRT_ERR_ARG, "rtrt_multi_attach_comm: bad communicator / rank / world");
    NcclApi& N = nccl_api();
    if (!N.ok) return c.fail(RTRT_ERR_NCCL, N.error.c_str());
    Multi* m = new Multi(); m->rank = rank; m->world = world; m->comm = reinterpret_cast<ncclComm_t>(nccl_comm); m->ownsComm = false;
    int rc = attach(c, m);
    if (rc) delete m;
    return rc;
}

int rtrt_create_multi(const rtrt_config* cfg, int device, const void* unique_id, int rank, int world, rtrt_ctx** out) {
    int rc = rtrt_create(cfg, device, out);
    if (rc) return rc;
    rc = rtrt_multi_init(*out, unique_id, rank, world);
    if (rc) { rtrt_destroy(*out); *out = nullptr; }
    return rc;
}

int rtrt_multi_link_local(rtrt_ctx** ctxs, int n) {
    if (!ctxs || n < 1) return RTRT_ERR_ARG;
    for (int i = 0; i < n; ++i) if (!ctxs[i] || ctxs[i]->c.multi) return RTRT_ERR_STATE;
    LocalGroup* g = new LocalGroup(); g->n = n; g->refs = n;
    for (int i = 0; i < n; ++i) g->members.push_back(&ctxs[i]->c);
    for (int i = 0; i < n; ++i) {
        Multi* m = new Multi(); m->rank = i; m->world = n; m->local = g;
        int rc = attach(ctxs[i]->c, m);
        if (rc) {                                               // undo: the i contexts attached so far leave the group, the last one frees it
            delete m;
            g->refs = i;
            if (i == 0) delete g;
            for (int j = 0; j < i; ++j) multi_release(ctxs[j]->c);
            return rc;
        }
    }
    return RTRT_OK;
}

int rtrt_multi_set_mode(rtrt_ctx* ctx, int mode, int strip_height) {
    if (!ctx) return RTRT_ERR_ARG;
    Context& c = ctx->c;
    Multi* m = multi_of(c);
    if (!m) return c.fail(RTRT_ERR_STATE, "rtrt_multi_set_mode: not a multi-GPU context");
    if (mode != RTRT_MULTI_ROWS && mode != RTRT_MULTI_SAMPLES) return c.fail(RTRT_ERR_ARG, "rtrt_multi_set_mode: unknown mode");
    if (strip_height < 1 || strip_height > 1024) return c.fail(RTRT_ERR_ARG, "rtrt_multi_set_mode: strip_height out of range");
    if (m->mode != mode || m->stripH != strip_height) { RTRT_CUDA(c, cudaSetDevice(c.device)); RTRT_CUDA(c, cudaStreamSynchronize(c.stream)); free_plans(*m); }
    m->mode = mode; m->stripH = strip_height;
    apply_partition(c, *m);
    return RTRT_OK;
}

int rtrt_multi_get_band(rtrt_ctx* ctx, int32_t* row_begin, int32_t* row_end, uint64_t* exchanges, uint64_t* bytes_sent) {
    if (!ctx) return RTRT_ERR_ARG;
    Context& c = ctx->c;
    Multi* m = multi_of(c);
    if (!m) return c.fail(RTRT_ERR_STATE, "rtrt_multi_get_band: not a multi-GPU context");
    std::vector<int> bb; band_bounds(c.height, m->world, bb);
    if (row_begin) *row_begin = bb[m->rank];
    if (row_end) *row_end = bb[m->rank + 1];
    if (exchanges) *exchanges = m->exchanges;
    if (bytes_sent) *bytes_sent = m->bytesSent;
    return RTRT_OK;
}

int rtrt_multi_plan_rows(int32_t height, int32_t world, int32_t strip_height, int32_t rank, int32_t peer, int32_t kind, int32_t halo, int32_t* rows_out, int32_t cap) {
    if (height < 1 || world < 1 || strip_height < 1 || rank < 0 || rank >= world || peer < 0 || peer >= world || kind < 0 || kind > 2 || halo < 0) return RTRT_ERR_ARG;
    if ((height + 15) / 16 < world) return RTRT_ERR_CAPACITY;
    std::vector<int> rows;
    plan_rows(height, world, strip_height, rank, peer, kind, halo, rows);
    if (rows_out) for (int i = 0; i < (int)rows.size() && i < cap; ++i) rows_out[i] = rows[i];
    return (int)rows.size();
}

// rank 0 passes the host arrays (pinned memory makes the upload asynchronous), the others NULL: upload on rank 0, broadcast over NVLink
int rtrt_update_vertices_multi(rtrt_ctx* ctx, const float* pos, const float* nrm, uint32_t V) {
    if (!ctx) return RTRT_ERR_ARG;
    Context& c = ctx->c;
    Multi* m = multi_of(c);
    if (!m) return c.fail(RTRT_ERR_STATE, "rtrt_update_vertices_multi: not a multi-GPU context");
    if (!c.dPos || !c.ownsMesh || V != c.numVerts) return c.fail(RTRT_ERR_STATE, "rtrt_update_vertices_multi: no library-owned mesh of that size");
    RTRT_CUDA(c, cudaSetDevice(c.device));
    if (m->rank == 0) {
        if (!pos) return c.fail(RTRT_ERR_ARG, "rtrt_update_vertices_multi: rank 0 must pass the positions");
        int rc = rtrt_update_vertices(ctx, pos, nrm, V);
        if (rc) return rc;
        rc = wait_geometry(c);                      // main stream waits for the upload
        if (rc) return rc;
    }
    int rc = broadcast0(c, *m, c.dPos, (size_t)3 * V * sizeof(float));
    if (rc) return rc;
    rc = broadcast0(c, *m, c.dNrm, (size_t)3 * V * sizeof(float));
    if (rc) return rc;
    c.bvhValid = false;
    return RTRT_OK;
}

int rtrt_render_frame_multi(rtrt_ctx* ctx) {
    if (!ctx) return RTRT_ERR_ARG;
    Context& c = ctx->c;
    Multi* m = multi_of(c);
    if (!m) return c.fail(RTRT_ERR_STATE, "rtrt_render_frame_multi: not a multi-GPU context (rtrt_create_multi / rtrt_multi_link_local)");
    RTRT_CUDA(c, cudaSetDevice(c.device));
    if (m->world == 1) return rtrt_render_frame(ctx);
    int rc;
    if (m->mode == RTRT_MULTI_SAMPLES) {
        // every rank: the whole frame with its own sample index; then the mean of the denoised frames
        const int base = c.frameNum;
        c.frameNum = base + m->rank;
        if ((rc = rtrt_build_bvh(ctx))) return rc;
        if ((rc = rtrt_pathtrace(ctx))) return rc;
        if ((rc = rtrt_denoise(ctx))) return rc;
        c.frameNum = base + m->world;
        const size_t px = (size_t)c.width * c.height;
        { rc = ensure_stage(c, *m, px * 8 * (m->local ? (size_t)m->world : 1)); if (rc) return rc; }
        // NCCL: reduce in a half4 copy held in recvStage; in-process: all_reduce stages through send / recv itself, so use the tail of... a separate image
        ushort4* img = reinterpret_cast<ushort4*>(c.dColorTmp);            // free between frames (the denoiser's ping-pong partner)
        k_rgb_to_stage<<<(unsigned)((px + 255) / 256), 256, 0, c.stream>>>(reinterpret_cast<const ushort4*>(c.dColor), img, px);
        if ((rc = all_reduce(c, *m, img, px * 8, true))) return rc;
        k_stage_to_rgb<<<(unsigned)((px + 255) / 256), 256, 0, c.stream>>>(reinterpret_cast<ushort4*>(c.dColor), img, px);
        c.launches += 2;
        RTRT_CUDA(c, cudaGetLastError());
        return RTRT_OK;
    }
    // ---- rows: strips for the path tracer, bands + halo exchange for the denoiser ----
    apply_partition(c, *m);
    if ((rc = rtrt_build_bvh(ctx))) return rc;
    if ((rc = rtrt_pathtrace(ctx))) return rc;
    {
        XBufs g{}; g.n = 5;
        void* ptrs[5] = {c.dColor, c.dNormal, c.dAlbedo, c.dMotion, c.dDepth}; const int bpp[5] = {8, 8, 8, 4, 2};
        for (int j = 0; j < 5; ++j) { g.ptr[j] = reinterpret_cast<unsigned char*>(ptrs[j]); g.rowBytes[j] = c.width * bpp[j]; }
        if ((rc = exchange(c, *m, X_GBUF, 0, g))) return rc;
    }
    if ((rc = rtrt_denoise(ctx))) return rc;                                // halo exchanges happen inside (Context::haloFn)
    {
        const bool temporal = c.passSettings.enableTemporalDenoising != 0, tf2 = c.passSettings.enableTemporalDenoising2 != 0;
        XBufs f{}; f.n = 0;
        f.ptr[f.n] = reinterpret_cast<unsigned char*>(c.dColor); f.rowBytes[f.n++] = c.width * 8;
        if (temporal) { f.ptr[f.n] = reinterpret_cast<unsigned char*>(c.dAccum); f.rowBytes[f.n++] = c.width * 8; }
        if ((rc = exchange(c, *m, X_FRAME, 0, f))) return rc;
        // the history colour image equals the final colour (k_dn_final stores both): complete it from the gathered frame
        if (tf2) RTRT_CUDA(c, cudaMemcpyAsync(c.dHistoryColor, c.dColor, (size_t)c.width * c.height * 8, cudaMemcpyDeviceToDevice, c.stream));
    }
    c.frameNum++;
    return RTRT_OK;
}

// BASELINE config 5: spp samples accumulated in fp32, partitioned BY SAMPLE (rank r renders samples r, r + N, ... of the whole frame), one
// all-reduce (sum) of the accumulation image at the end.  On return RTRT_BUF_ACCUM_RADIANCE holds the sum over all spp samples on every rank.
int rtrt_render_progressive_multi(rtrt_ctx* ctx, int32_t spp, int32_t first_frame) {
    if (!ctx) return RTRT_ERR_ARG;
    Context& c = ctx->c;
    Multi* m = multi_of(c);
    if (!m) return c.fail(RTRT_ERR_STATE, "rtrt_render_progressive_multi: not a multi-GPU context");
    if (spp < 1) return c.fail(RTRT_ERR_ARG, "rtrt_render_progressive_multi: spp must be >= 1");
    RTRT_CUDA(c, cudaSetDevice(c.device));
    const int rb = c.rowBegin, re = c.rowEnd, sh = c.stripH, ss = c.stripShards, sk = c.stripShard;
    c.rowBegin = 0; c.rowEnd = c.height; c.stripH = 0; c.stripShards = 1; c.stripShard = 0;
    int rc = rtrt_build_bvh(ctx);
    int mine = 0;
    for (int s = m->rank; !rc && s < spp; s += m->world) {
        c.frameNum = first_frame + s;
        rc = rtrt_pathtrace(ctx);
        if (!rc) rc = rtrt_accumulate(ctx, mine == 0);
        ++mine;
    }
    if (!rc && mine == 0) {                                                // more ranks than samples: contribute zeros
        c.frameNum = first_frame; rc = rtrt_pathtrace(ctx);
        if (!rc) rc = rtrt_accumulate(ctx, 1);
        if (!rc && cudaMemsetAsync(c.dAccumRadiance, 0, (size_t)c.width * c.height * 16, c.stream) != cudaSuccess) rc = RTRT_ERR_CUDA;
    }
    if (!rc) rc = all_reduce(c, *m, c.dAccumRadiance, (size_t)c.width * c.height * 16, false);
    c.rowBegin = rb; c.rowEnd = re; c.stripH = sh; c.stripShards = ss; c.stripShard = sk;
    return rc;
}

}  // extern "C"
