// "Fast" traversal layout (option wide_bvh): an 8-wide BVH with child boxes quantised to 8 bits inside the parent box and
// 48-byte position-only triangles, derived on the device from the reference-layout (parity) arrays after every build.
//
// Why: B200 has no RT cores.  The reference's layout (src/bvhNode.cuh:5-13: 64-byte two-child nodes, src/geometry.h:52-97: 128-byte
// triangles, ONE triangle per leaf; loop src/traverse.h:132-250) costs ~70 dependent node fetches + 3.3 leaf fetches per incoherent
// ray on the 1 M-triangle bench scene, and that chain — not bandwidth — bounds the path tracer (profiles/r1j_ncu_summary.md).
// An 8-wide node answers three binary levels with ONE cache line, leaves hold up to 3 triangles, and the whole node array of the bench
// scene is ~9 MB (vs 64 MB), i.e. it lives in L2.
//
// Layout (node = 128-byte stride, first 80 bytes used — one L1/L2 line per visit; after Ylitie, Karras, Laine 2017,
// "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs"):
//   row 0   float px, py, pz      origin = min corner of the node box
//           u8 ex, ey, ez         IEEE exponent bytes of the per-axis grid step (step = 2^(e-127), 255 steps >= extent)
//           u8 imask              bit s = child slot s is an inner node
//   row 1   u32 childBase         index of the first inner child; inner children are consecutive in slot order
//           u32 triBase           index (in triangles) of the node's first leaf triangle in the wide triangle array
//           u8 meta[8]            per slot: 0 empty | inner: 0x20 | (24 + s) | leaf of k <= 3 triangles: (2^k - 1) << 5 | first triangle (0..23)
//   row 2-4 u8 qlo[3][8], qhi[3][8]   child boxes in grid steps: rows are {lox, loy}, {loz, hix}, {hiy, hiz}, 8 bytes each
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

namespace rtrt {

constexpr int kWideStride = 8;                   // float4 rows per node (128 B)
constexpr int kWideRegion = 512;                 // node slots reserved per 1024-triangle batch (bound: see wide_build.cu)
constexpr int kWideMaxLeaf = 3;                  // triangles per leaf slot
constexpr int kWideStackSmem = 16;               // stack entries per thread in shared memory (8 B each: the reference stack's footprint) ...
constexpr int kWideStackSpill = 48;              // ... and in local memory behind them (short shared stack with spill)

struct WideRay {
    Float3 org;
    float idx, idy, idz;                         // 1 / dir (safe_divide, src/traverse.cuh:88)
    float aox, aoy, aoz;                         // |org| * |1/dir| * 2^-20: absolute-coordinate part of the box-test margin
    uint32_t octinv4;                            // (7 - octant) replicated into 4 bytes
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
__device__ __forceinline__ float byte_to_float(uint32_t v, int i) {   // byte i of v as float (exact)
    return (float)((v >> (8 * i)) & 0xFFu);
}

// Box tests of the four children whose quantised planes are in (lox, loy, loz, hix, hiy, hiz): each word holds one plane byte per child.
// t = q * a + b per plane (one FMA; a = step / dir, b = (origin - ray origin) / dir), near planes biased by -delta, far planes by +delta.
__device__ __forceinline__ uint32_t wide_test4(uint32_t meta4, uint32_t octinv4, uint32_t nx, uint32_t ny, uint32_t nz, uint32_t fx, uint32_t fy, uint32_t fz,
                                               float ax, float ay, float az, float bnx, float bny, float bnz, float bfx, float bfy, float bfz, float tmax) {
    const uint32_t isInner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
    const uint32_t innerMask4 = sign_extend_s8x4(isInner4 << 3);
    const uint32_t bitIndex4 = (meta4 ^ (octinv4 & innerMask4)) & 0x1F1F1F1Fu;
    const uint32_t childBits4 = (meta4 >> 5) & 0x07070707u;
    uint32_t hit = 0u;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float tnx = fmaf(byte_to_float(nx, i), ax, bnx), tny = fmaf(byte_to_float(ny, i), ay, bny), tnz = fmaf(byte_to_float(nz, i), az, bnz);
        const float tfx = fmaf(byte_to_float(fx, i), ax, bfx), tfy = fmaf(byte_to_float(fy, i), ay, bfy), tfz = fmaf(byte_to_float(fz, i), az, bfz);
        const float cmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
        const float cmax = fminf(fminf(tfx, tfy), fminf(tfz, tmax));
        if (cmin <= cmax) hit |= ((childBits4 >> (8 * i)) & 0xFFu) << ((bitIndex4 >> (8 * i)) & 0xFFu);
    }
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

// One traversal step: visit ONE inner node (if the current group has any), test ALL triangles it reported, pop when the group is
// used up.  Returns true when the ray is finished.
template <bool COUNT>
__device__ __forceinline__ bool wide_step(const WideView& wv, const WideRay& w, WideState& s, int anyHit, const WideStack& st) {
    {   // the current group always has a pending inner child here: it is the root group, or it was popped (groups are only pushed
        // with children left), or the node visited in the previous step reported inner hits
        const uint32_t hits = s.ngroup.y;
        const uint32_t bit = 31u - (uint32_t)__clz(hits);
        s.ngroup.y &= ~(1u << bit);
        if (s.ngroup.y > 0x00FFFFFFu) wide_push(st, s, s.ngroup);
        const uint32_t slot = (bit - 24u) ^ (w.octinv4 & 0xFFu);
        const uint32_t rel = __popc(hits & ~(0xFFFFFFFFu << slot) & 0xFFu);
        const float4* nd = wv.nodes + (size_t)(s.ngroup.x + rel) * kWideStride;
        const float4 n0 = __ldg(nd), n1 = __ldg(nd + 1), n2 = __ldg(nd + 2), n3 = __ldg(nd + 3), n4 = __ldg(nd + 4);
        if (COUNT) s.nodeVisits++;
        const uint32_t eim = __float_as_uint(n0.w);
        const float ax = __uint_as_float((eim & 0xFFu) << 23) * w.idx;
        const float ay = __uint_as_float(((eim >> 8) & 0xFFu) << 23) * w.idy;
        const float az = __uint_as_float(((eim >> 16) & 0xFFu) << 23) * w.idz;
        const float bx = (n0.x - w.org.x) * w.idx, by = (n0.y - w.org.y) * w.idy, bz = (n0.z - w.org.z) * w.idz;
        // error margin of t = q a + b against exact arithmetic on the stored planes (3 eps |b| + 2 eps |q a|, eps = 2^-24) and of the
        // watertight test's own translation v - org (eps |coordinate| / |dir|): 2^-20 (|b| + 256 |a| + (|org| + |origin|) / |dir|)
        const float m = 9.5367431640625e-07f;
        const float dx = m * (fabsf(bx) + 256.0f * fabsf(ax) + fabsf(n0.x) * fabsf(w.idx)) + w.aox;
        const float dy = m * (fabsf(by) + 256.0f * fabsf(ay) + fabsf(n0.y) * fabsf(w.idy)) + w.aoy;
        const float dz = m * (fabsf(bz) + 256.0f * fabsf(az) + fabsf(n0.z) * fabsf(w.idz)) + w.aoz;
        const float bnx = bx - dx, bny = by - dy, bnz = bz - dz, bfx = bx + dx, bfy = by + dy, bfz = bz + dz;
        // near plane of an axis = lo plane when the direction is positive, hi plane otherwise
        const bool px = w.idx >= 0.0f, py = w.idy >= 0.0f, pz = w.idz >= 0.0f;
        const uint32_t lox0 = __float_as_uint(n2.x), lox1 = __float_as_uint(n2.y), loy0 = __float_as_uint(n2.z), loy1 = __float_as_uint(n2.w);
        const uint32_t loz0 = __float_as_uint(n3.x), loz1 = __float_as_uint(n3.y), hix0 = __float_as_uint(n3.z), hix1 = __float_as_uint(n3.w);
        const uint32_t hiy0 = __float_as_uint(n4.x), hiy1 = __float_as_uint(n4.y), hiz0 = __float_as_uint(n4.z), hiz1 = __float_as_uint(n4.w);
        uint32_t hitmask = wide_test4(__float_as_uint(n1.z), w.octinv4, px ? lox0 : hix0, py ? loy0 : hiy0, pz ? loz0 : hiz0,
                                      px ? hix0 : lox0, py ? hiy0 : loy0, pz ? hiz0 : loz0, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, s.tSlack);
        hitmask |= wide_test4(__float_as_uint(n1.w), w.octinv4, px ? lox1 : hix1, py ? loy1 : hiy1, pz ? loz1 : hiz1,
                              px ? hix1 : lox1, py ? hiy1 : loy1, pz ? hiz1 : loz1, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, s.tSlack);
        s.ngroup = make_uint2(__float_as_uint(n1.x), (hitmask & 0xFF000000u) | (eim >> 24));
        s.tgroup = make_uint2(__float_as_uint(n1.y), hitmask & 0x00FFFFFFu);
    }
    if (s.tgroup.y != 0u) {
        const RayPrep rp = wide_tri_prep(w);
        do {
            const uint32_t bit = 31u - (uint32_t)__clz(s.tgroup.y);
            s.tgroup.y &= ~(1u << bit);
            const float4* row = wv.tris + (size_t)(s.tgroup.x + bit) * 3;
            const float4 q1 = __ldg(row), q2 = __ldg(row + 1), q3 = __ldg(row + 2);
            if (COUNT) s.leafTests++;
            float tt, uu, vv;
            if (ray_triangle(rp, q1, q2, q3, s.t, tt, uu, vv)) {
                if (tt < s.t) {
                    s.t = tt; s.tSlack = tt * 1.000001f; s.tri = __float_as_int(q1.w); s.ua = uu; s.va = vv;
                    if (anyHit) return true;
                }
            }
        } while (s.tgroup.y != 0u);
    }
    if (s.ngroup.y <= 0x00FFFFFFu) {
        if (s.sp == 0) return true;
        s.ngroup = wide_pop(st, s);
    }
    return false;
}

// ---- one interface over both layouts, so the traversal kernels (k_traverse, k_pt_primary, k_pt_trace) exist once ----
// `raw` = the thread's slice of a shared buffer of 16 x 8 bytes per thread: (link, tNear) columns of the reference stack, or the
// first 16 (base, mask) entries of the wide stack.
struct ParityTrav {
    static constexpr bool kWide = false;
    RayPrep rp; TravState s;
    uint32_t* stL; float* stT; int stride, cap, maxLoop; uint32_t numTri; AABB sceneBox;
    __device__ __forceinline__ void setup(const SceneView& sc, uint2* raw, int tid, int threads) {
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
    WideRay w; WideState s; WideStack st;
    uint2 spill[kWideStackSpill];
    __device__ __forceinline__ void setup(const SceneView& sc, uint2* raw, int tid, int threads) {
        st.smem = raw + tid; st.stride = threads; st.spill = spill;
        begin(sc, Float3{0, 0, 0}, Float3{0, 1, 0});
    }
    __device__ __forceinline__ void begin(const SceneView& sc, const Float3& o, const Float3& d) { w = wide_prepare_ray(o, d); wide_init(s, sc.wide.root); }
    template <bool COUNT>
    __device__ __forceinline__ bool step(const SceneView& sc, int anyHit) { return wide_step<COUNT>(sc.wide, w, s, anyHit, st); }
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
