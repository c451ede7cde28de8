// "Fast" traversal layout (option wide_bvh): an 8-wide BVH with fp16 child boxes relative to the parent box and
// 48-byte position-only triangles, derived on the device from the reference-layout (parity) arrays after every build.
//
// Why: B200 has no RT cores.  The reference's layout (src/bvhNode.cuh:5-13: 64-byte two-child nodes, src/geometry.h:52-97: 128-byte
// triangles, ONE triangle per leaf; loop src/traverse.h:132-250) costs ~70 dependent node fetches + 3.3 leaf fetches per incoherent
// ray on the 1 M-triangle bench scene, and that chain — not bandwidth — bounds the path tracer (profiles/r1j_ncu_summary.md).
// An 8-wide node answers three binary levels with ONE cache line, a leaf slot holds one triangle (kWideMaxLeaf), and the whole node array of the bench
// scene is 28 MB (+ 50 MB of 48-byte triangles; the reference layout's traversal reads 64 + 134 MB), i.e. it lives in the 126 MB L2.
//
// Layout (node = 128 bytes = one L1/L2 line per visit; the slot / octant scheme follows Ylitie, Karras, Laine 2017, "Efficient
// Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs"):
//   row 0   float px, py, pz      origin = min corner of the node box
//           u8 ex, ey, ez         IEEE exponent bytes of the per-axis scale 2^(e-127) >= extent (child planes are stored in units of it)
//           u8 imask              bit s = child slot s is an inner node
//   row 1   u32 childBase         index of the first inner child; inner children are consecutive in slot order
//           u32 triBase           index (in triangles) of the node's first leaf triangle in the wide triangle array
//           u8 meta[8]            per slot: 0 empty | inner: 0x20 | (24 + s) | leaf of k <= 3 triangles: (2^k - 1) << 5 | first triangle (0..23)
//   row 2-7 half lox[8], loy[8], loz[8], hix[8], hiy[8], hiz[8]   child boxes relative to the origin in units of the scale, in [0, 1]:
//           fp16, lo planes rounded down, hi planes rounded up.  Round 2 first used 8-bit grid coordinates (Ylitie et al.): decoding 48 bytes
//           per visit (PRMT / I2F) kept the ALU pipe 57 % busy and bound the kernel (profiles/r2d); a half converts to float on the FMA pipe
//           (HADD2.F32 with a half selector, no extraction), the grid is 8x finer, and the node still is one cache line.
// Slots are assigned so that (slot ^ octant) approximates front-to-back order for every ray octant (no per-visit sort, no
// distances on the stack).  Triangle = 3 x float4 {v.xyz, w}: row0.w carries the ORIGINAL triangle index (batch * 1024 + local id),
// so hit records, normals and materials keep using the reference-layout arrays.
//
// Parity contract of the mode: the triangle test is the reference's RayTriangleWatertight (traverse.cuh: ray_triangle, same
// operand bits), the box test is conservative (explicit error margins below), so every ray finds the same closest (tri, t, u, v)
// as an UNCAPPED reference traversal; it differs from the reference only for the rays on which the reference itself drops
// pushes on its full 16-entry stack or hits its 1024-iteration cap, and on exact ties in t between two triangles (first found
// wins in both, the visiting order differs).  tests/test_wide_bvh_gpu.py counts all three classes.
#pragma once
#include "traverse.cuh"
#include <cuda_fp16.h>

namespace rtrt {

constexpr int kWideStride = 8;                   // float4 rows per node (128 B)
constexpr int kWideRegion = 512;                 // node slots reserved per 1024-triangle batch (bound: see wide_build.cu)
#ifndef RTRT_WIDE_MAX_LEAF
#define RTRT_WIDE_MAX_LEAF 1      // measured on the bench frame, 3 -> 2 -> 1: triangle tests per bounce ray 6.9 -> 5.1 -> 3.4, node visits 23.0 -> 23.5 -> 24.1,
                                  // path trace 2.29 -> 2.20 -> 2.12 ms, collapse 0.22 -> 0.24 -> 0.29 ms: the ~100-instruction triangle test runs with few lanes
#endif
constexpr int kWideMaxLeaf = RTRT_WIDE_MAX_LEAF;  // triangles per leaf slot (<= 3: the meta byte has three triangle bits)
constexpr int kWideStackSmem = 16;               // stack entries per thread in shared memory (8 B each: the reference stack's footprint) ...
constexpr int kWideStackSpill = 48;              // ... and in local memory behind them (short shared stack with spill)

struct WideRay {
    Float3 org;
    float idx, idy, idz;                         // 1 / dir (safe_divide, src/traverse.cuh:88)
    float aox, aoy, aoz;                         // |org| * |1/dir| * 2^-20: absolute-coordinate part of the box-test margin
    uint32_t octinv4;                            // (7 - octant) replicated into 4 bytes
    int rnx, rny, rnz;                           // node row of the NEAR plane per axis: lo row (2, 3, 4) for a non-negative reciprocal, else hi row (5, 6, 7)
    int kz; float Sx, Sy, Sz;                    // watertight test constants (RayPrep subset)
};

struct WideState {
    uint2 ngroup;                                // x = child base index, y = hits of inner children [31:24] | imask [7:0]
    uint2 tgroup;                                // x = triangle base index, y = hit triangles [23:0]
    int sp;                                      // stack entries in use
    float t, tSlack; int tri; float ua, va;
    uint32_t nodeVisits, leafTests, overflow;
};

__device__ __forceinline__ WideRay wide_prepare_ray(const Float3& org, const Float3& dir) {
    WideRay w;
    w.org = org;
    w.idx = safe_divide(1.0f, dir.x); w.idy = safe_divide(1.0f, dir.y); w.idz = safe_divide(1.0f, dir.z);
    const float m = 9.5367431640625e-07f;        // 2^-20
    w.aox = fabsf(org.x) * fabsf(w.idx) * m; w.aoy = fabsf(org.y) * fabsf(w.idy) * m; w.aoz = fabsf(org.z) * fabsf(w.idz) * m;
    const uint32_t oct = (dir.x < 0.0f ? 4u : 0u) | (dir.y < 0.0f ? 2u : 0u) | (dir.z < 0.0f ? 1u : 0u);
    w.octinv4 = (7u - oct) * 0x01010101u;
    w.rnx = w.idx >= 0.0f ? 2 : 5; w.rny = w.idy >= 0.0f ? 3 : 6; w.rnz = w.idz >= 0.0f ? 4 : 7;
    int kz = max_dim3(fabsf(dir.x), fabsf(dir.y), fabsf(dir.z));
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    w.kz = kz;
    w.Sx = comp(dir, kx) / comp(dir, kz);
    w.Sy = comp(dir, ky) / comp(dir, kz);
    w.Sz = 1.0f / comp(dir, kz);
    return w;
}

// the watertight test wants a RayPrep; only these members are read by ray_triangle (the rest is dead and never materialised)
__device__ __forceinline__ RayPrep wide_tri_prep(const WideRay& w) {
    RayPrep r;
    r.org = w.org; r.kz = w.kz; r.Sx = w.Sx; r.Sy = w.Sy; r.Sz = w.Sz;
    return r;
}

__device__ __forceinline__ void wide_init(WideState& s, uint32_t root) {
    s.ngroup = make_uint2(root, 0x80000000u);    // "slot 7 ^ octinv of a group whose imask is 0" -> node index root + 0
    s.tgroup = make_uint2(0u, 0u);
    s.sp = 0;
    s.t = kRayMax; s.tSlack = kRayMax; s.tri = -1; s.ua = 0.0f; s.va = 0.0f;
    s.nodeVisits = 0; s.leafTests = 0; s.overflow = 0;
}

__device__ __forceinline__ uint32_t sign_extend_s8x4(uint32_t x) {   // 0x80 in a byte -> 0xFF, else 0x00
    uint32_t r;
    asm("prmt.b32 %0, %1, 0, 0xba98;" : "=r"(r) : "r"(x));            // replicate-sign mode
    return r;
}
// child planes: 8 halves per 16-byte row; half I of the row as float (one HADD2.F32 on the FMA pipe, the half is selected in the instruction)
template <int I>
__device__ __forceinline__ float row_half(const uint4& r) {
    const uint32_t wd = (I >> 1) == 0 ? r.x : ((I >> 1) == 1 ? r.y : ((I >> 1) == 2 ? r.z : r.w));
    const __half2 h = *reinterpret_cast<const __half2*>(&wd);
    return (I & 1) ? __high2float(h) : __low2float(h);
}

// Box tests of the four children LO..LO+3 (meta4 = their meta bytes): t = l A + b per plane, l = the stored plane in [0, 1], A = scale / dir,
// b = (origin - ray origin) / dir; near planes carry -delta, far planes +delta (wide_step).
template <int LO>
__device__ __forceinline__ uint32_t wide_test4(uint32_t meta4, uint32_t octinv4, const uint4& nx, const uint4& ny, const uint4& nz, const uint4& fx, const uint4& fy,
                                               const uint4& fz, float Ax, float Ay, float Az, float cnx, float cny, float cnz, float cfx, float cfy, float cfz, float tmax) {
    const uint32_t isInner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
    const uint32_t innerMask4 = sign_extend_s8x4(isInner4 << 3);
    const uint32_t bitIndex4 = (meta4 ^ (octinv4 & innerMask4)) & 0x1F1F1F1Fu;
    const uint32_t childBits4 = (meta4 >> 5) & 0x07070707u;
    uint32_t hit = 0u;
#define RTRT_WIDE_CHILD(I)                                                                                                                              \
    {                                                                                                                                                   \
        const float tnx = fmaf(row_half<LO + I>(nx), Ax, cnx), tny = fmaf(row_half<LO + I>(ny), Ay, cny), tnz = fmaf(row_half<LO + I>(nz), Az, cnz);   \
        const float tfx = fmaf(row_half<LO + I>(fx), Ax, cfx), tfy = fmaf(row_half<LO + I>(fy), Ay, cfy), tfz = fmaf(row_half<LO + I>(fz), Az, cfz);   \
        const float cmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));                                                                                    \
        const float cmax = fminf(fminf(tfx, tfy), fminf(tfz, tmax));                                                                                    \
        if (cmin <= cmax) hit |= ((childBits4 >> (8 * I)) & 0xFFu) << ((bitIndex4 >> (8 * I)) & 0xFFu);                                                 \
    }
    RTRT_WIDE_CHILD(0) RTRT_WIDE_CHILD(1) RTRT_WIDE_CHILD(2) RTRT_WIDE_CHILD(3)
#undef RTRT_WIDE_CHILD
    return hit;
}

// Stack of (base, mask) pairs: the first kWideStackSmem entries of a thread live in shared memory (stride = threads of the CTA),
// deeper ones in a per-thread local array (only ever touched by rays that need them).
struct WideStack {
    uint2* smem; int stride; uint2* spill;
};
__device__ __forceinline__ void wide_push(const WideStack& st, WideState& s, uint2 v) {
    if (s.sp < kWideStackSmem) st.smem[s.sp * st.stride] = v;
    else if (s.sp < kWideStackSmem + kWideStackSpill) st.spill[s.sp - kWideStackSmem] = v;
    else { s.overflow++; return; }                                  // counted, never silent (tests assert 0)
    ++s.sp;
}
__device__ __forceinline__ uint2 wide_pop(const WideStack& st, WideState& s) {
    --s.sp;
    return s.sp < kWideStackSmem ? st.smem[s.sp * st.stride] : st.spill[s.sp - kWideStackSmem];
}

// One traversal step of a whole warp (every lane calls it, `active` = the lane has a ray in flight):
//   fetch      a lane whose node group is used up takes the next group off its stack;
//   node       every lane with a pending inner child visits ONE node: 8 box tests -> hit inner children + hit triangles;
//   triangles  POSTPONED (Ylitie et al. 2017, section 5.2): the hit triangles of a node wait in the lane's pending group (an older
//              pending group goes onto the stack) and are tested, one per lane and step, only when the warp votes that enough lanes
//              have one pending (>= RTRT_WIDE_TRI_NUM / RTRT_WIDE_TRI_DEN of the lanes in flight), or when some lane has nothing else
//              left to do.  The first cut tested all triangles of a node right after it: ~45 % of the kernel's issue slots ran the
//              ~80-instruction triangle test with ~3 of 32 lanes (profiles/r2b: 15.6 lanes per instruction).
// The closest hit does not depend on the test order (up to exact ties in t), so (tri, t, u, v) are unchanged.
// Returns true when this lane's ray is finished.
#ifndef RTRT_WIDE_TRI_NUM
#define RTRT_WIDE_TRI_NUM 3
#endif
#ifndef RTRT_WIDE_TRI_DEN
#define RTRT_WIDE_TRI_DEN 8
#endif
#ifndef RTRT_WIDE_TRI_MUST
#define RTRT_WIDE_TRI_MUST 1
#endif
__device__ __forceinline__ uint2 wide_peek(const WideStack& st, const WideState& s) {
    const int i = s.sp - 1;
    return i < kWideStackSmem ? st.smem[i * st.stride] : st.spill[i - kWideStackSmem];
}
// stack entries: node group (hit mask of inner children in y[31:24]) or triangle group (y[31:24] == 0, hit triangles in y[23:0])
__device__ __forceinline__ void wide_fetch(const WideStack& st, WideState& s) {
    if (s.ngroup.y <= 0x00FFFFFFu && s.sp > 0) {
        const uint2 e = wide_peek(st, s);
        if (e.y > 0x00FFFFFFu) { s.ngroup = e; --s.sp; }                     // more nodes first; a pending triangle group keeps waiting
        else if (s.tgroup.y == 0u) { s.tgroup = e; --s.sp; }
    }
}

template <bool COUNT>
__device__ __forceinline__ bool wide_step(const WideView& wv, const WideRay& w, WideState& s, int anyHit, const WideStack& st, bool active) {
    const unsigned full = 0xffffffffu;
    if (active && s.ngroup.y > 0x00FFFFFFu) {
        const uint32_t hits = s.ngroup.y;
        const uint32_t bit = 31u - (uint32_t)__clz(hits);
        s.ngroup.y &= ~(1u << bit);
        if (s.ngroup.y > 0x00FFFFFFu) wide_push(st, s, s.ngroup);
        const uint32_t slot = (bit - 24u) ^ (w.octinv4 & 0xFFu);
        const uint32_t rel = __popc(hits & ~(0xFFFFFFFFu << slot) & 0xFFu);
        const float4* nd = wv.nodes + (size_t)(s.ngroup.x + rel) * kWideStride;
        const uint4* nu = reinterpret_cast<const uint4*>(nd);
        const float4 n0 = __ldg(nd), n1 = __ldg(nd + 1);
        // near / far plane rows are picked by ADDRESS (per-ray constants), not by selects on the loaded data.  (Fetching the node as four
        // 32-byte loads — LDG.E.256, sm_100 — halves the L1 tag look-ups, 113 M -> 64 M sectors per launch, but then the rows must be selected
        // in registers: measured 0.926 -> 0.950 ms for the bounce round and 0.67 -> 0.75 ms for the primary rays, profiles/r2f; not kept.)
        const uint4 nx = __ldg(nu + w.rnx), ny = __ldg(nu + w.rny), nz = __ldg(nu + w.rnz);
        const uint4 fx = __ldg(nu + (7 - w.rnx)), fy = __ldg(nu + (9 - w.rny)), fz = __ldg(nu + (11 - w.rnz));
        if (COUNT) s.nodeVisits++;
        const uint32_t eim = __float_as_uint(n0.w);
        const float Ax = __uint_as_float((eim & 0xFFu) << 23) * w.idx;
        const float Ay = __uint_as_float(((eim >> 8) & 0xFFu) << 23) * w.idy;
        const float Az = __uint_as_float(((eim >> 16) & 0xFFu) << 23) * w.idz;
        const float bx = (n0.x - w.org.x) * w.idx, by = (n0.y - w.org.y) * w.idy, bz = (n0.z - w.org.z) * w.idz;
        // Error margin delta of t = l A + b against exact arithmetic on the stored planes (l exact, eps = 2^-24): A one rounding (eps |A|),
        // b two roundings plus the cancellation in origin - org (2 eps |b| + eps (|org| + |origin|) / |dir|), the fma eps |t| <= eps (|A| + |b|),
        // and the watertight test's own translation v - org (eps |coordinate| / |dir|): all below 2^-20 (|A| + |b| + (|org| + |origin|) / |dir|).
        const float m = 9.5367431640625e-07f;
        const float dx = m * (fabsf(bx) + fabsf(Ax) + fabsf(n0.x) * fabsf(w.idx)) + w.aox;
        const float dy = m * (fabsf(by) + fabsf(Ay) + fabsf(n0.y) * fabsf(w.idy)) + w.aoy;
        const float dz = m * (fabsf(bz) + fabsf(Az) + fabsf(n0.z) * fabsf(w.idz)) + w.aoz;
        const float bnx = bx - dx, bny = by - dy, bnz = bz - dz, bfx = bx + dx, bfy = by + dy, bfz = bz + dz;
        uint32_t hitmask = wide_test4<0>(__float_as_uint(n1.z), w.octinv4, nx, ny, nz, fx, fy, fz, Ax, Ay, Az, bnx, bny, bnz, bfx, bfy, bfz, s.tSlack);
        hitmask |= wide_test4<4>(__float_as_uint(n1.w), w.octinv4, nx, ny, nz, fx, fy, fz, Ax, Ay, Az, bnx, bny, bnz, bfx, bfy, bfz, s.tSlack);
        s.ngroup = make_uint2(__float_as_uint(n1.x), (hitmask & 0xFF000000u) | (eim >> 24));
        if (hitmask & 0x00FFFFFFu) {
            if (s.tgroup.y != 0u) wide_push(st, s, s.tgroup);                 // the older pending group waits on the stack
            s.tgroup = make_uint2(__float_as_uint(n1.y), hitmask & 0x00FFFFFFu);
        }
    }
    if (active) wide_fetch(st, s);
    // ---- triangle phase, warp-voted ----
    const bool want = active && s.tgroup.y != 0u;
    const bool must = want && s.ngroup.y <= 0x00FFFFFFu;                      // (after the fetch) nothing but triangles left to do
    const unsigned wm = __ballot_sync(full, want), mm = __ballot_sync(full, must), am = __ballot_sync(full, active);
    bool finished = false;
#if RTRT_WIDE_TRI_MUST
    const bool force = mm != 0u;                                              // a lane with nothing but triangles left forces the phase
#else
    const bool force = __ballot_sync(full, active && s.ngroup.y > 0x00FFFFFFu) == 0u;   // ... only when NO lane has node work left: lanes that can only
                                                                              // test triangles wait for company instead of running the test with a few lanes
#endif
    if (wm != 0u && (force || __popc(wm) * RTRT_WIDE_TRI_DEN >= __popc(am) * RTRT_WIDE_TRI_NUM)) {
        if (want) {
            const RayPrep rp = wide_tri_prep(w);
            const uint32_t bit = 31u - (uint32_t)__clz(s.tgroup.y);
            s.tgroup.y &= ~(1u << bit);
            const float4* row = wv.tris + (size_t)(s.tgroup.x + bit) * 3;
            const float4 q1 = __ldg(row), q2 = __ldg(row + 1), q3 = __ldg(row + 2);
            if (COUNT) s.leafTests++;
            float tt, uu, vv;
            if (ray_triangle(rp, q1, q2, q3, s.t, tt, uu, vv)) {
                if (tt < s.t) {
                    s.t = tt; s.tSlack = tt * 1.000001f; s.tri = __float_as_int(q1.w); s.ua = uu; s.va = vv;
                    if (anyHit) finished = true;
                }
            }
            if (!finished) wide_fetch(st, s);
        }
    }
    if (active && s.ngroup.y <= 0x00FFFFFFu && s.tgroup.y == 0u && s.sp == 0) finished = true;
    return finished;
}

// ---- one interface over both layouts, so the traversal kernels (k_traverse, k_pt_primary, k_pt_trace) exist once ----
// `raw` = the thread's slice of a shared buffer of 16 x 8 bytes per thread: (link, tNear) columns of the reference stack, or the
// first 16 (base, mask) entries of the wide stack.
struct ParityTrav {
    static constexpr bool kWide = false;
    RayPrep rp; TravState s;
    uint32_t* stL; float* stT; int stride, cap, maxLoop; uint32_t numTri; AABB sceneBox;
    static constexpr int kSpill = 1;                     // no spill storage: the reference drops pushes on a full stack
    __device__ __forceinline__ void setup(const SceneView& sc, uint2* raw, int tid, int threads, uint2*) {
        uint32_t* base = reinterpret_cast<uint32_t*>(raw);
        stL = base + tid; stT = reinterpret_cast<float*>(base + 16 * threads) + tid; stride = threads;
        cap = sc.stackSize < 16 ? sc.stackSize : 16; maxLoop = sc.maxLoop; numTri = sc.numBatches * kBatchSize;
        const float4 a = __ldg(sc.tlasRows), b = __ldg(sc.tlasRows + 1), c = __ldg(sc.tlasRows + 2);   // merged box of TLAS node 0 (src/traverse.h:127-129)
        sceneBox.max = Float3{sel_max(a.x, c.x), sel_max(a.y, c.y), sel_max(a.z, c.z)};
        sceneBox.min = Float3{sel_min(b.x, a.w), sel_min(b.y, b.w), sel_min(b.z, c.w)};
        begin(sc, Float3{0, 0, 0}, Float3{0, 1, 0});
    }
    __device__ __forceinline__ void begin(const SceneView&, const Float3& o, const Float3& d) { rp = prepare_ray(o, d, sceneBox); trav_init(s); }
    template <bool COUNT>
    __device__ __forceinline__ bool step(const SceneView& sc, int anyHit) { return trav_step<COUNT>(sc, rp, s, anyHit, cap, numTri, maxLoop, stL, stT, stride); }
    // called by every lane of the warp (the wide layout votes inside); true = this lane's ray is finished
    template <bool COUNT>
    __device__ __forceinline__ bool step_warp(const SceneView& sc, int anyHit, bool active) { return active ? step<COUNT>(sc, anyHit) : false; }
    __device__ __forceinline__ float t() const { return s.t; }
    __device__ __forceinline__ int tri() const { return s.tri; }
    __device__ __forceinline__ float u() const { return s.u; }
    __device__ __forceinline__ float v() const { return s.v; }
    __device__ __forceinline__ float ua() const { return s.ua; }
    __device__ __forceinline__ float va() const { return s.va; }
    __device__ __forceinline__ uint32_t iterations() const { return (uint32_t)s.it; }
    __device__ __forceinline__ uint32_t nodeVisits() const { return s.nodeVisits; }
    __device__ __forceinline__ uint32_t leafTests() const { return s.leafTests; }
};

struct WideTrav {
    static constexpr bool kWide = true;
    // The spill array is NOT a member: an array member that is indexed dynamically puts the whole object (ray constants and traversal
    // state included) into local memory.  The kernel declares `uint2 spill[kWideStackSpill]` next to its Trav and passes it to setup().
    static constexpr int kSpill = kWideStackSpill;
    WideRay w; WideState s; WideStack st;
    __device__ __forceinline__ void setup(const SceneView& sc, uint2* raw, int tid, int threads, uint2* spill) {
        st.smem = raw + tid; st.stride = threads; st.spill = spill;
        begin(sc, Float3{0, 0, 0}, Float3{0, 1, 0});
    }
    __device__ __forceinline__ void begin(const SceneView& sc, const Float3& o, const Float3& d) { w = wide_prepare_ray(o, d); wide_init(s, sc.wide.root); }
    template <bool COUNT>
    __device__ __forceinline__ bool step_warp(const SceneView& sc, int anyHit, bool active) { return wide_step<COUNT>(sc.wide, w, s, anyHit, st, active); }
    __device__ __forceinline__ float t() const { return s.t; }
    __device__ __forceinline__ int tri() const { return s.tri; }
    __device__ __forceinline__ float u() const { return s.ua; }      // the closest hit's barycentrics (the reference's u, v differ from them only on exact ties)
    __device__ __forceinline__ float v() const { return s.va; }
    __device__ __forceinline__ float ua() const { return s.ua; }
    __device__ __forceinline__ float va() const { return s.va; }
    __device__ __forceinline__ uint32_t iterations() const { return s.nodeVisits + s.leafTests; }
    __device__ __forceinline__ uint32_t nodeVisits() const { return s.nodeVisits; }
    __device__ __forceinline__ uint32_t leafTests() const { return s.leafTests; }
};

}  // namespace rtrt
