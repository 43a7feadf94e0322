// avr_graph_file.cpp — reader / writer for the fork's graph text files: Graph::WriteToStream /
// ReadFromStream (src/graph/graph.cpp:284-416), FreeGraph::WriteToStream / ReadFromStream (:644-655),
// Graph::WriteToDisk (:419-433); field layout documented in the reference's files/format.txt.
// Pure host code behind the C ABI (include/avr.h); byte-compatible with what `graph_maker` writes:
// SEP = ' ' after every field (also before the newline), floats as `ostream << float` prints them
// (precision 6, %g), "True"/"False" flags.
#include <cerrno>
#include <cstdio>
#include <new>
#include <stdexcept>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/avr.h"

namespace avr {
avr_status set_err(avr_ctx *ctx, avr_status code, const std::string &msg);  // avr_lib.cu
}
using namespace avr;

namespace {

struct Tokens {
    std::vector<char> buf;
    char *p = nullptr, *end = nullptr;
    bool load(const char *path) {
        FILE *f = std::fopen(path, "rb");
        if (!f) return false;
        std::fseek(f, 0, SEEK_END);
        long n = std::ftell(f);
        std::fseek(f, 0, SEEK_SET);
        if (n < 0) { std::fclose(f); return false; }  // not seekable
        buf.resize(size_t(n) + 1);
        size_t got = std::fread(buf.data(), 1, size_t(n), f);
        std::fclose(f);
        buf[got] = 0;
        p = buf.data();
        end = p + got;
        return true;
    }
    void skip() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p; }
    bool word(std::string *w) {  // istream >> std::string
        skip();
        if (p >= end) return false;
        char *q = p;
        while (q < end && !(*q == ' ' || *q == '\n' || *q == '\t' || *q == '\r')) ++q;
        w->assign(p, q);
        p = q;
        return true;
    }
    bool i32(int32_t *v) {
        skip();
        if (p >= end) return false;
        char *q = nullptr;
        errno = 0;
        long x = std::strtol(p, &q, 10);
        if (q == p || errno == ERANGE || x < INT32_MIN || x > INT32_MAX) return false;
        *v = int32_t(x);
        p = q;
        return true;
    }
    bool f32(float *v) {
        skip();
        if (p >= end) return false;
        char *q = nullptr;
        float x = std::strtof(p, &q);
        if (q == p) return false;
        *v = x;
        p = q;
        return true;
    }
};

template <typename T>
T *take(std::vector<T> &v) {  // hand a vector's contents to the C struct (malloc'd copy)
    if (v.empty()) return nullptr;
    T *p = (T *)std::malloc(v.size() * sizeof(T));
    std::memcpy(p, v.data(), v.size() * sizeof(T));
    return p;
}

avr_status fail(avr_status code, const std::string &msg) { return set_err(nullptr, code, msg); }

}  // namespace

extern "C" void avr_graph_file_free(avr_graph_file *g) {
    if (!g) return;
    void *ptrs[] = {g->vertex_id, g->vertex_pos, g->vertex_type, g->vertex_light, g->vertex_samples, g->vertex_range,
                    g->vertex_coors, g->edge_id, g->edge_from, g->edge_to, g->edge_samples, g->path_id, g->path_offset,
                    g->path_vertices};
    for (void *p : ptrs) std::free(p);
    std::free(g);
}

extern "C" avr_status avr_graph_file_read(const char *path, avr_graph_file **out) {
    try {
        if (!path || !out) return fail(AVR_ERR_BAD_ARG, "avr_graph_file_read: NULL argument");
        *out = nullptr;
        Tokens t;
        if (!t.load(path)) return fail(AVR_ERR_BAD_ARG, std::string("avr_graph_file_read: cannot open ") + path);
        std::string desc, name, flag[5];
        float radius = -1;
        // FreeGraph::ReadFromStream: in >> description >> name >> vertexRadius
        if (!t.word(&desc) || !t.word(&name) || !t.f32(&radius))
            return fail(AVR_ERR_BAD_ARG, "graph file: truncated header");
        if (name != "free")
            return fail(AVR_ERR_UNSUPPORTED, "graph file: only \"free\" graphs are on the path (got \"" + name + "\")");
        for (int i = 0; i < 5; ++i)
            if (!t.word(&flag[i])) return fail(AVR_ERR_BAD_ARG, "graph file: truncated flags line");
        int32_t hdr[6];
        for (int i = 0; i < 6; ++i)
            if (!t.i32(&hdr[i])) return fail(AVR_ERR_BAD_ARG, "graph file: truncated counts line");
        if (hdr[3] < 0 || hdr[4] < 0 || hdr[5] < 0) return fail(AVR_ERR_BAD_ARG, "graph file: negative count");
        bool coors = flag[0] == "True", samples = flag[1] == "True", types = flag[2] == "True", light = flag[3] == "True",
             range = flag[4] == "True";
        size_t nv = size_t(hdr[3]), ne = size_t(hdr[4]), np = size_t(hdr[5]);
        // a record is at least "id x y z\n" / "id from to\n" / "id n\n": counts a file of this size cannot hold are a
        // corrupt (or hostile) header, not a reason to allocate gigabytes
        size_t file_bytes = size_t(t.end - t.buf.data());
        if (nv > file_bytes / 8 || ne > file_bytes / 6 || np > file_bytes / 4)
            return fail(AVR_ERR_BAD_ARG, "graph file: the header counts exceed what the file can hold");
        std::vector<int32_t> vid(nv), vtype, vsmp, vco, eid(ne), efrom(ne), eto(ne), esmp, pid(np), pverts;
        std::vector<float> vpos(3 * nv), vlight, vrange;
        std::vector<int64_t> poff(np + 1, 0);
        if (types) vtype.resize(nv);
        if (light) vlight.resize(nv);
        if (samples) { vsmp.resize(nv); esmp.resize(ne); }
        if (range) vrange.resize(nv);
        if (coors) vco.resize(3 * nv);
        bool ok = true;
        for (size_t i = 0; i < nv && ok; ++i) {
            ok = t.i32(&vid[i]) && t.f32(&vpos[3 * i]) && t.f32(&vpos[3 * i + 1]) && t.f32(&vpos[3 * i + 2]);
            // ReadVertexData order: type, lightScalar, samples, renderSearchRange (graph.cpp:249-267)
            if (ok && types) ok = t.i32(&vtype[i]);
            if (ok && light) ok = t.f32(&vlight[i]);
            if (ok && samples) ok = t.i32(&vsmp[i]);
            if (ok && range) ok = t.f32(&vrange[i]);
            if (ok && coors) ok = t.i32(&vco[3 * i]) && t.i32(&vco[3 * i + 1]) && t.i32(&vco[3 * i + 2]);
        }
        for (size_t i = 0; i < ne && ok; ++i) {
            ok = t.i32(&eid[i]) && t.i32(&efrom[i]) && t.i32(&eto[i]);
            if (ok && samples) ok = t.i32(&esmp[i]);
        }
        for (size_t i = 0; i < np && ok; ++i) {
            int32_t len = 0;
            ok = t.i32(&pid[i]) && t.i32(&len) && len >= 0;
            for (int32_t j = 0; j < len && ok; ++j) {
                int32_t v;
                ok = t.i32(&v);
                if (ok) pverts.push_back(v);
            }
            poff[i + 1] = int64_t(pverts.size());
        }
        if (!ok) return fail(AVR_ERR_BAD_ARG, "graph file: truncated or malformed record");
        avr_graph_file *g = (avr_graph_file *)std::calloc(1, sizeof(avr_graph_file));
        std::snprintf(g->description, sizeof(g->description), "%s", desc.c_str());
        g->vertex_radius = radius;
        g->use_coors = coors; g->use_samples = samples; g->use_ray_vertex_types = types; g->use_lighting = light;
        g->use_render_search_range = range;
        g->cur_vertex_id = hdr[0]; g->cur_edge_id = hdr[1]; g->cur_path_id = hdr[2];
        g->n_vertices = hdr[3]; g->n_edges = hdr[4]; g->n_paths = hdr[5];
        g->vertex_id = take(vid); g->vertex_pos = take(vpos); g->vertex_type = take(vtype); g->vertex_light = take(vlight);
        g->vertex_samples = take(vsmp); g->vertex_range = take(vrange); g->vertex_coors = take(vco);
        g->edge_id = take(eid); g->edge_from = take(efrom); g->edge_to = take(eto); g->edge_samples = take(esmp);
        g->path_id = take(pid); g->path_offset = take(poff); g->path_vertices = take(pverts);
        *out = g;
        return AVR_OK;
    } catch (const std::bad_alloc &) {
        return fail(AVR_ERR_OOM, "avr_graph_file_read: out of memory");
    } catch (const std::exception &e) {
        return fail(AVR_ERR_BAD_ARG, std::string("avr_graph_file_read: ") + e.what());
    }
}

extern "C" avr_status avr_graph_file_write(const char *path, const avr_graph_file *g) {
    if (!path || !g) return fail(AVR_ERR_BAD_ARG, "avr_graph_file_write: NULL argument");
    size_t nv = size_t(g->n_vertices), ne = size_t(g->n_edges), np = size_t(g->n_paths);
    if ((nv && (!g->vertex_id || !g->vertex_pos)) || (ne && (!g->edge_id || !g->edge_from || !g->edge_to)) ||
        (np && (!g->path_id || !g->path_offset)))
        return fail(AVR_ERR_BAD_ARG, "avr_graph_file_write: missing vertex / edge / path arrays");
    if ((g->use_ray_vertex_types && nv && !g->vertex_type) || (g->use_lighting && nv && !g->vertex_light) ||
        (g->use_samples && ((nv && !g->vertex_samples) || (ne && !g->edge_samples))) ||
        (g->use_render_search_range && nv && !g->vertex_range) || (g->use_coors && nv && !g->vertex_coors))
        // the reference: ErrorExit("Graph stream options or flags not set") / optional::value() throws
        return fail(AVR_ERR_BAD_ARG, "avr_graph_file_write: a flag is set but its array is NULL");
    FILE *f = std::fopen(path, "wb");
    if (!f) return fail(AVR_ERR_BAD_ARG, std::string("avr_graph_file_write: cannot open ") + path);
    auto tf = [](int32_t b) { return b ? "True" : "False"; };
    std::fprintf(f, "%s\n", g->description);
    std::fprintf(f, "free %g\n", g->vertex_radius);
    std::fprintf(f, "%s %s %s %s %s\n", tf(g->use_coors), tf(g->use_samples), tf(g->use_ray_vertex_types),
                 tf(g->use_lighting), tf(g->use_render_search_range));
    std::fprintf(f, "%d %d %d %zu %zu %zu\n", g->cur_vertex_id, g->cur_edge_id, g->cur_path_id, nv, ne, np);
    for (size_t i = 0; i < nv; ++i) {
        std::fprintf(f, "%d %g %g %g ", g->vertex_id[i], g->vertex_pos[3 * i], g->vertex_pos[3 * i + 1],
                     g->vertex_pos[3 * i + 2]);
        if (g->use_ray_vertex_types) std::fprintf(f, "%d ", g->vertex_type[i]);
        if (g->use_lighting) std::fprintf(f, "%g ", g->vertex_light[i]);
        if (g->use_samples) std::fprintf(f, "%d ", g->vertex_samples[i]);
        if (g->use_render_search_range) std::fprintf(f, "%g ", g->vertex_range[i]);
        if (g->use_coors)
            std::fprintf(f, "%d %d %d ", g->vertex_coors[3 * i], g->vertex_coors[3 * i + 1], g->vertex_coors[3 * i + 2]);
        std::fputc('\n', f);
    }
    for (size_t i = 0; i < ne; ++i) {
        std::fprintf(f, "%d %d %d ", g->edge_id[i], g->edge_from[i], g->edge_to[i]);
        if (g->use_samples) std::fprintf(f, "%d ", g->edge_samples[i]);
        std::fputc('\n', f);
    }
    for (size_t i = 0; i < np; ++i) {
        int64_t a = g->path_offset[i], b = g->path_offset[i + 1];
        std::fprintf(f, "%d %lld ", g->path_id[i], (long long)(b - a));
        for (int64_t j = a; j < b; ++j) std::fprintf(f, "%d ", g->path_vertices[j]);
        std::fputc('\n', f);
    }
    bool bad = std::ferror(f) != 0;
    bad = (std::fclose(f) != 0) || bad;
    if (bad) return fail(AVR_ERR_BAD_ARG, std::string("avr_graph_file_write: write error on ") + path);
    return AVR_OK;
}

// ---- binary CSR sidecar (include/avr.h: avr_graph_sidecar) -------------------------------------------
namespace {

struct Fnv {
    uint64_t h = 0xcbf29ce484222325ull;
    void add(const void *p, size_t n) {
        const unsigned char *b = (const unsigned char *)p;
        for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 0x100000001b3ull; }
    }
};
struct Writer {
    FILE *f;
    Fnv fnv;
    bool ok = true;
    void put(const void *p, size_t n) {
        if (!n) return;
        fnv.add(p, n);
        ok = ok && std::fwrite(p, 1, n, f) == n;
    }
};
struct Reader {
    FILE *f;
    Fnv fnv;
    bool get(void *p, size_t n) {
        if (!n) return true;
        if (std::fread(p, 1, n, f) != n) return false;
        fnv.add(p, n);
        return true;
    }
};
enum { SC_POS = 1, SC_LIGHT = 2, SC_RANGE = 4 };

// NULL if rowptr / col describe a CSR matrix over n_vertices rows and columns
const char *csr_defect(const avr_graph_sidecar *s) {
    size_t n = size_t(s->n_vertices), nnz = size_t(s->nnz);
    if (s->rowptr[0] != 0 || s->rowptr[n] != s->nnz) return "rowptr does not span [0, nnz]";
    for (size_t i = 0; i < n; ++i)
        if (s->rowptr[i + 1] < s->rowptr[i]) return "rowptr not monotone";
    for (size_t i = 0; i < nnz; ++i)
        if (s->col[i] < 0 || s->col[i] >= s->n_vertices) return "column index outside [0, n)";
    return nullptr;
}

}  // namespace

extern "C" void avr_graph_sidecar_free(avr_graph_sidecar *s) {
    if (!s) return;
    void *ptrs[] = {s->rowptr, s->col, s->val, s->vertex_pos, s->light, s->search_range};
    for (void *p : ptrs) std::free(p);
    std::free(s);
}

extern "C" avr_status avr_graph_sidecar_write(const char *path, const avr_graph_sidecar *s) {
    if (!path || !s) return fail(AVR_ERR_BAD_ARG, "avr_graph_sidecar_write: NULL argument");
    if (s->n_vertices < 1 || s->nnz < 0 || !s->rowptr || (s->nnz && (!s->col || !s->val)))
        return fail(AVR_ERR_BAD_ARG, "avr_graph_sidecar_write: need n_vertices >= 1, rowptr, and col / val for nnz > 0");
    if (const char *why = csr_defect(s))
        return fail(AVR_ERR_BAD_ARG, std::string("avr_graph_sidecar_write: ") + why);
    FILE *f = std::fopen(path, "wb");
    if (!f) return fail(AVR_ERR_BAD_ARG, std::string("avr_graph_sidecar_write: cannot open ") + path);
    Writer w{f};
    size_t n = size_t(s->n_vertices), nnz = size_t(s->nnz);
    uint32_t version = 1, flags = (s->vertex_pos ? SC_POS : 0) | (s->light ? SC_LIGHT : 0) | (s->search_range ? SC_RANGE : 0);
    w.put("AVRG", 4);
    w.put(&version, 4);
    w.put(&s->n_vertices, 4);
    w.put(&flags, 4);
    w.put(&s->nnz, 8);
    w.put(s->rowptr, (n + 1) * 8);
    w.put(s->col, nnz * 4);
    w.put(s->val, nnz * 4);
    if (s->vertex_pos) w.put(s->vertex_pos, 3 * n * 4);
    if (s->light) w.put(s->light, n * 4);
    if (s->search_range) w.put(s->search_range, n * 4);
    uint64_t sum = w.fnv.h;
    bool ok = w.ok && std::fwrite(&sum, 1, 8, f) == 8;
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) return fail(AVR_ERR_BAD_ARG, std::string("avr_graph_sidecar_write: write error on ") + path);
    return AVR_OK;
}

extern "C" avr_status avr_graph_sidecar_read(const char *path, avr_graph_sidecar **out) {
    if (!path || !out) return fail(AVR_ERR_BAD_ARG, "avr_graph_sidecar_read: NULL argument");
    *out = nullptr;
    FILE *f = std::fopen(path, "rb");
    if (!f) return fail(AVR_ERR_BAD_ARG, std::string("avr_graph_sidecar_read: cannot open ") + path);
    std::fseek(f, 0, SEEK_END);
    long file_bytes = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    Reader r{f};
    char magic[4];
    uint32_t version = 0, flags = 0;
    int32_t n = 0;
    int64_t nnz = 0;
    avr_graph_sidecar *s = nullptr;
    auto bail = [&](const char *msg) {
        std::fclose(f);
        avr_graph_sidecar_free(s);
        return fail(AVR_ERR_BAD_ARG, std::string("graph sidecar: ") + msg);
    };
    if (!r.get(magic, 4) || std::memcmp(magic, "AVRG", 4) != 0) return bail("not an AVRG file");
    if (!r.get(&version, 4) || version != 1) return bail("unsupported version");
    if (!r.get(&n, 4) || !r.get(&flags, 4) || !r.get(&nnz, 8) || n < 1 || nnz < 0 || (flags & ~7u))
        return bail("bad header");
    // the header fixes the file size: check it before allocating from untrusted counts
    unsigned long long per_vertex = 8ull + ((flags & SC_POS) ? 12ull : 0ull) + ((flags & SC_LIGHT) ? 4ull : 0ull) +
                                    ((flags & SC_RANGE) ? 4ull : 0ull);
    unsigned long long want = 24ull + 8ull + per_vertex * (unsigned long long)n + 8ull * (unsigned long long)nnz + 8ull;
    if ((unsigned long long)file_bytes != want) return bail("file size does not match its header (truncated?)");
    s = (avr_graph_sidecar *)std::calloc(1, sizeof(avr_graph_sidecar));
    s->n_vertices = n;
    s->nnz = nnz;
    size_t N = size_t(n), NNZ = size_t(nnz);
    s->rowptr = (int64_t *)std::malloc((N + 1) * 8);
    s->col = (int32_t *)std::malloc(NNZ ? NNZ * 4 : 4);
    s->val = (float *)std::malloc(NNZ ? NNZ * 4 : 4);
    if (flags & SC_POS) s->vertex_pos = (float *)std::malloc(3 * N * 4);
    if (flags & SC_LIGHT) s->light = (float *)std::malloc(N * 4);
    if (flags & SC_RANGE) s->search_range = (float *)std::malloc(N * 4);
    if (!s->rowptr || !s->col || !s->val || ((flags & SC_POS) && !s->vertex_pos) || ((flags & SC_LIGHT) && !s->light) ||
        ((flags & SC_RANGE) && !s->search_range))
        return bail("out of memory");
    bool ok = r.get(s->rowptr, (N + 1) * 8) && r.get(s->col, NNZ * 4) && r.get(s->val, NNZ * 4);
    if (ok && s->vertex_pos) ok = r.get(s->vertex_pos, 3 * N * 4);
    if (ok && s->light) ok = r.get(s->light, N * 4);
    if (ok && s->search_range) ok = r.get(s->search_range, N * 4);
    uint64_t sum = 0;
    if (!ok || std::fread(&sum, 1, 8, f) != 8) return bail("truncated");
    if (sum != r.fnv.h) return bail("checksum mismatch");
    if (const char *why = csr_defect(s)) return bail(why);
    std::fclose(f);
    *out = s;
    return AVR_OK;
}
