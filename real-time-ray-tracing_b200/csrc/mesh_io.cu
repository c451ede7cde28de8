// Mesh ingest formats either side of the path, host code:
//   .bin   uint32 triCount; Triangle[triCount]   reader  LoadTrianglesFromFile      src/init.cu:28-50
//                                                 writer  tool/meshProcessor.cpp:204-208
//          current layout 128 B/triangle (v1..v4, n1..n4 as float3 + pad, src/geometry.h:52-110); the legacy layout of the shipped
//          resources/models/test.bin is 64 B/triangle (v1..v4 only) — told apart by the file size, since both share the header
//   .obj   Wavefront OBJ.  The reference reads meshes through Assimp (src/fileUtils.cu:16-80, aiProcess_Triangulate), an
//          un-vendored dependency; this is a direct reader of the subset its assets use: v / vn / f with v, v/vt, v//vn, v/vt/vn
//          corners, negative (relative) indices, polygons split as a fan from the first corner (what Assimp's triangulation emits
//          for triangles and convex quads).  Triangles come out in file order = Assimp's node/mesh/face order for single-object files.
//   pre-sort  tool/meshProcessor.cpp:146-200 sorts the triangles by the 60-bit Morton code of their centroid before writing; here
//          the order comes from the device pre-sort of bvh_presort.cu (stable: equal codes keep the input order, where the tool's
//          std::sort leaves them unspecified).
#include "context.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace rtrt;

namespace {

struct Soup { std::vector<float> pos, nrm; bool hasNormals = true; uint32_t tris() const { return (uint32_t)(pos.size() / 9); } };

int upload_soup(rtrt_ctx* ctx, const Soup& s) {
    const uint32_t n = s.tris();
    std::vector<uint32_t> idx((size_t)3 * n);
    for (uint32_t i = 0; i < 3 * n; ++i) idx[i] = i;
    return rtrt_set_mesh(ctx, s.pos.data(), s.hasNormals ? s.nrm.data() : nullptr, idx.data(), 3 * n, n);
}

// one face corner "v", "v/vt", "v//vn", "v/vt/vn"; returns false on a malformed token
bool parse_corner(const char*& p, long& v, long& vn, bool& hasVn) {
    char* end = nullptr;
    v = strtol(p, &end, 10);
    if (end == p) return false;
    p = end; hasVn = false; vn = 0;
    if (*p == '/') {
        ++p;
        if (*p != '/') { strtol(p, &end, 10); p = end; }              // vt (ignored)
        if (*p == '/') {
            ++p;
            vn = strtol(p, &end, 10);
            if (end == p) return false;
            p = end; hasVn = true;
        }
    }
    return true;
}

int read_bin(const char* path, Soup& s, std::string& err) {
    FILE* f = fopen(path, "rb");
    if (!f) { err = "cannot open file"; return RTRT_ERR_IO; }
    uint32_t n = 0;
    if (fread(&n, sizeof(n), 1, f) != 1 || n == 0) { fclose(f); err = "bad header"; return RTRT_ERR_IO; }
    fseek(f, 0, SEEK_END);
    const long size = ftell(f);
    fseek(f, 4, SEEK_SET);
    const bool legacy = size == 4 + 64L * (long)n;
    if (!legacy && size < 4 + 128L * (long)n) { fclose(f); err = "truncated file"; return RTRT_ERR_IO; }
    const size_t rec = legacy ? 64 : 128;
    std::vector<unsigned char> raw(rec * n);
    if (fread(raw.data(), rec, n, f) != n) { fclose(f); err = "truncated file"; return RTRT_ERR_IO; }
    fclose(f);
    // un-indexed triangle soup (3 vertices per triangle); the legacy layout has no normals (generated on upload)
    s.pos.resize((size_t)9 * n); s.hasNormals = !legacy;
    if (!legacy) s.nrm.resize((size_t)9 * n);
    for (uint32_t i = 0; i < n; ++i)
        for (int k = 0; k < 3; ++k) {
            std::memcpy(&s.pos[(size_t)9 * i + 3 * k], &raw[rec * i + 16 * k], 12);
            if (!legacy) std::memcpy(&s.nrm[(size_t)9 * i + 3 * k], &raw[rec * i + 64 + 16 * k], 12);
        }
    return RTRT_OK;
}

int read_obj(const char* path, Soup& s, std::string& err) {
    FILE* f = fopen(path, "rb");
    if (!f) { err = "cannot open file"; return RTRT_ERR_IO; }
    std::vector<float> v, vn;
    std::string line;
    char buf[4096];
    long lineNo = 0;
    auto bad = [&](const char* what) { fclose(f); err = std::string(what) + " at line " + std::to_string(lineNo); return (int)RTRT_ERR_IO; };
    while (fgets(buf, sizeof(buf), f)) {
        line = buf;
        while (!line.empty() && line.back() != '\n' && fgets(buf, sizeof(buf), f)) line += buf;    // lines longer than the buffer
        ++lineNo;
        const char* p = line.c_str();
        while (*p == ' ' || *p == '\t') ++p;
        if (p[0] == 'v' && (p[1] == ' ' || p[1] == '\t')) {
            float x[3]; char* end = nullptr; p += 2;
            for (int k = 0; k < 3; ++k) { x[k] = strtof(p, &end); if (end == p) return bad("malformed vertex"); p = end; }
            v.insert(v.end(), x, x + 3);
        } else if (p[0] == 'v' && p[1] == 'n' && (p[2] == ' ' || p[2] == '\t')) {
            float x[3]; char* end = nullptr; p += 3;
            for (int k = 0; k < 3; ++k) { x[k] = strtof(p, &end); if (end == p) return bad("malformed normal"); p = end; }
            vn.insert(vn.end(), x, x + 3);
        } else if (p[0] == 'f' && (p[1] == ' ' || p[1] == '\t')) {
            p += 2;
            std::vector<long> cv, cn;
            bool allVn = true;
            for (;;) {
                while (*p == ' ' || *p == '\t') ++p;
                if (*p == '\0' || *p == '\n' || *p == '\r' || *p == '#') break;
                long a, b; bool h;
                if (!parse_corner(p, a, b, h)) return bad("malformed face");
                const long nv = (long)(v.size() / 3), nn = (long)(vn.size() / 3);
                a = a < 0 ? nv + a : a - 1;                                         // relative / 1-based
                if (a < 0 || a >= nv) return bad("vertex index out of range");
                if (h) { b = b < 0 ? nn + b : b - 1; if (b < 0 || b >= nn) return bad("normal index out of range"); }
                allVn = allVn && h;
                cv.push_back(a); cn.push_back(h ? b : -1);
            }
            if (cv.size() < 3) return bad("face with fewer than 3 corners");
            for (size_t k = 1; k + 1 < cv.size(); ++k) {                            // fan from the first corner
                const size_t tri[3] = {0, k, k + 1};
                for (size_t t : tri) {
                    s.pos.insert(s.pos.end(), &v[3 * cv[t]], &v[3 * cv[t]] + 3);
                    if (allVn) s.nrm.insert(s.nrm.end(), &vn[3 * cn[t]], &vn[3 * cn[t]] + 3);
                    else { const float z[3] = {0, 0, 0}; s.nrm.insert(s.nrm.end(), z, z + 3); }
                }
            }
            s.hasNormals = s.hasNormals && allVn;                                   // mesh->HasNormals(): all or nothing
        }
    }
    fclose(f);
    if (s.tris() == 0) { err = "no faces in file"; return RTRT_ERR_IO; }
    return RTRT_OK;
}

bool has_ext(const char* path, const char* ext) {
    const size_t n = strlen(path), m = strlen(ext);
    if (n < m) return false;
    for (size_t i = 0; i < m; ++i) { char a = path[n - m + i]; if (a >= 'A' && a <= 'Z') a += 32; if (a != ext[i]) return false; }
    return true;
}

}  // namespace

extern "C" {

// Context-free (pure host) reader, so the formats can be checked without a GPU.  `.obj` by extension, anything else = `.bin`.
int rtrt_mesh_file_read(const char* path, float** tri_pos, float** tri_nrm, uint32_t* num_tris, char* err, size_t err_len) {
    if (!path || !tri_pos || !tri_nrm || !num_tris) return RTRT_ERR_ARG;
    *tri_pos = *tri_nrm = nullptr; *num_tris = 0;
    Soup s; std::string e;
    const int rc = has_ext(path, ".obj") ? read_obj(path, s, e) : read_bin(path, s, e);
    if (rc) { if (err && err_len) snprintf(err, err_len, "%s", e.c_str()); return rc; }
    const size_t bytes = s.pos.size() * sizeof(float);
    *tri_pos = (float*)malloc(bytes);
    if (!*tri_pos) return RTRT_ERR_CAPACITY;
    std::memcpy(*tri_pos, s.pos.data(), bytes);
    if (s.hasNormals) {
        *tri_nrm = (float*)malloc(bytes);
        if (!*tri_nrm) { free(*tri_pos); *tri_pos = nullptr; return RTRT_ERR_CAPACITY; }
        std::memcpy(*tri_nrm, s.nrm.data(), bytes);
    }
    *num_tris = s.tris();
    return RTRT_OK;
}

void rtrt_mesh_file_free(void* p) { free(p); }

int rtrt_set_mesh_bin(rtrt_ctx* ctx, const char* path) {
    if (!ctx) return RTRT_ERR_ARG;
    Context& c = ctx->c;
    if (!path) return c.fail(RTRT_ERR_ARG, "rtrt_set_mesh_bin: null path");
    Soup s; std::string e;
    const int rc = read_bin(path, s, e);
    if (rc) return c.fail(rc, ("rtrt_set_mesh_bin: " + e).c_str());
    return upload_soup(ctx, s);
}

int rtrt_set_mesh_obj(rtrt_ctx* ctx, const char* path) {
    if (!ctx) return RTRT_ERR_ARG;
    Context& c = ctx->c;
    if (!path) return c.fail(RTRT_ERR_ARG, "rtrt_set_mesh_obj: null path");
    Soup s; std::string e;
    const int rc = read_obj(path, s, e);
    if (rc) return c.fail(rc, ("rtrt_set_mesh_obj: " + e).c_str());
    return upload_soup(ctx, s);
}

int rtrt_write_mesh_bin(rtrt_ctx* ctx, const char* path, int presort, int legacy64) {
    if (!ctx) return RTRT_ERR_ARG;
    Context& c = ctx->c;
    if (!path) return c.fail(RTRT_ERR_ARG, "rtrt_write_mesh_bin: null path");
    if (!c.dPos || !c.dIdx || c.numTris == 0) return c.fail(RTRT_ERR_STATE, "rtrt_write_mesh_bin: no mesh set");
    RTRT_CUDA(c, cudaSetDevice(c.device));
    const uint32_t N = c.numTris, V = c.numVerts;
    std::vector<float> pos((size_t)3 * V), nrm((size_t)3 * V);
    std::vector<uint32_t> idx((size_t)3 * N), perm;
    { int rc = wait_geometry(c); if (rc) return rc; }
    if (presort) {
        int rc = launch_presort(c);                                                  // k_ps_bounds -> k_ps_codes -> radix sort -> k_ps_gather
        if (rc) return rc;
        perm.resize(N);
        RTRT_CUDA(c, cudaMemcpyAsync(perm.data(), c.dTriPerm, (size_t)4 * N, cudaMemcpyDeviceToHost, c.stream));
    }
    RTRT_CUDA(c, cudaMemcpyAsync(pos.data(), c.dPos, (size_t)12 * V, cudaMemcpyDeviceToHost, c.stream));
    if (c.dNrm) RTRT_CUDA(c, cudaMemcpyAsync(nrm.data(), c.dNrm, (size_t)12 * V, cudaMemcpyDeviceToHost, c.stream));
    RTRT_CUDA(c, cudaMemcpyAsync(idx.data(), c.dIdx, (size_t)12 * N, cudaMemcpyDeviceToHost, c.stream));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    const size_t rec = legacy64 ? 64 : 128;
    std::vector<unsigned char> raw(rec * N, 0);                                      // pads / v4 / n4 zero (the reference leaves them uninitialised)
    for (uint32_t i = 0; i < N; ++i) {
        const uint32_t t = presort ? perm[i] : i;
        for (int k = 0; k < 3; ++k) {
            const uint32_t vi = idx[(size_t)3 * t + k];
            if (vi >= V) return c.fail(RTRT_ERR_STATE, "rtrt_write_mesh_bin: index out of range");
            std::memcpy(&raw[rec * i + 16 * k], &pos[(size_t)3 * vi], 12);
            if (!legacy64) std::memcpy(&raw[rec * i + 64 + 16 * k], &nrm[(size_t)3 * vi], 12);
        }
    }
    FILE* f = fopen(path, "wb");
    if (!f) return c.fail(RTRT_ERR_IO, "rtrt_write_mesh_bin: cannot open file for writing");
    const bool ok = fwrite(&N, 4, 1, f) == 1 && fwrite(raw.data(), rec, N, f) == N;
    if (fclose(f) != 0 || !ok) return c.fail(RTRT_ERR_IO, "rtrt_write_mesh_bin: write failed");
    return RTRT_OK;
}

}  // extern "C"
