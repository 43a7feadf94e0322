// avr_nvdb.cpp — ".nvdb" ingest: what NanoVDBMedium::Create does with its "filename"
// (src/pbrt/media.cpp:487-509 readGrid, :628-665 Create) before the constructor runs — find the named FloatGrid
// in the file, check that it is a fog volume — and the flattening of the grid over its index bounding box that the
// fork's own cmd/nanovdb2pbrt.cpp:100-110 does voxel by voxel through tree().getValue().  Pure host code.
//
// The reference reads the file with the NanoVDB headers it vendors (src/ext/openvdb @ 414bed84, branch
// feature/nanovdb: ABI / file-format major version 32).  Those headers are not part of this repository; the
// on-disk layout of that version is restated here (offsets below; oracle/ prints them from the vendored headers,
// tests/test_nvdb.py reads files the reference's own nanovdb::io::writeGrid wrote):
//   file    = segments; segment = Header{magic u64, version u32, gridCount u16, codec u16}
//             + gridCount x (MetaData[176] + name[nameSize]) + gridCount x grid buffer (fileSize bytes each;
//             codec 0 = raw, 1 = ZIP: u64 size + zlib stream, 2 = BLOSC: not supported here)
//   buffer  = GridData[672] TreeData[64] ... ; TreeData.mNodeOffset[4] = byte offsets from the tree of
//             {leaf, lower, upper, root} arrays
//   root    = RootData[64]{bbox, tableSize, background...} + tiles[32]{key u64, child i64 (offset from the root, 0 =
//             value tile), state u32, value f32}
//   upper   = 32^3 table at +8256 (8 bytes each: float value | i64 child offset from the node), child mask at
//             +4128, node size 270400 ; lower = 16^3 table at +1088, child mask at +544, node size 33856
//   leaf    = 8^3 floats at +96, node size 2144 ; voxel (x, y, z) at ((x & 7) << 6) | ((y & 7) << 3) | (z & 7)
// Instead of 10^8 getValue() calls the tree is walked once: leaves are copied, value tiles are filled.
#include <dlfcn.h>

#include <algorithm>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/avr.h"

namespace avr {
avr_status set_err(avr_ctx *ctx, avr_status code, const std::string &msg);
}

namespace {

constexpr uint64_t kMagic = 0x304244566f6e614eull;  // "NanoVDB0"
constexpr int kMajor = 32;
constexpr size_t kHeader = 16, kMeta = 176, kGridData = 672, kTreeData = 64, kRootData = 64, kRootTile = 32;
constexpr size_t kUpperSize = 270400, kUpperChildMask = 4128, kUpperTable = 8256;
constexpr size_t kLowerSize = 33856, kLowerChildMask = 544, kLowerTable = 1088;
constexpr size_t kLeafSize = 2144, kLeafValues = 96;
// GridData members
constexpr size_t kGdVersion = 16, kGdGridSize = 32, kGdMap = 296, kGdWorldBBox = 560, kGdGridClass = 632, kGdGridType = 636;
constexpr size_t kMapInvMatF = 36, kMapVecF = 72;
// MetaData members
constexpr size_t kMdGridSize = 0, kMdFileSize = 8, kMdGridType = 32, kMdGridClass = 36, kMdNameSize = 136, kMdCodec = 168;

template <typename T>
T rd(const unsigned char *p) {
    T v;
    std::memcpy(&v, p, sizeof(T));
    return v;
}

avr_status fail(avr_status code, const std::string &msg) { return avr::set_err(nullptr, code, msg); }

struct GridEntry {
    std::string name;
    uint64_t grid_size = 0, file_size = 0, data_pos = 0;
    uint32_t grid_type = 0, grid_class = 0;
    uint16_t codec = 0;
};

// every grid of every segment of the file
avr_status list_grids(FILE *f, const std::string &path, std::vector<GridEntry> *out) {
    std::fseek(f, 0, SEEK_END);
    long file_bytes = std::ftell(f);
    if (file_bytes < 0) return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": cannot seek");
    uint64_t pos = 0;
    while (pos + kHeader <= uint64_t(file_bytes)) {
        unsigned char h[kHeader];
        std::fseek(f, long(pos), SEEK_SET);
        if (std::fread(h, 1, kHeader, f) != kHeader) break;
        if (rd<uint64_t>(h) != kMagic) {
            if (pos == 0) return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": Magic number error: This is not a valid nvdb file");
            break;
        }
        uint32_t version = rd<uint32_t>(h + 8);
        if (int(version >> 21) != kMajor)
            return fail(AVR_ERR_UNSUPPORTED, "nanovdb: " + path + ": Incompatible file format: major version " +
                                                 std::to_string(version >> 21) + ", this reader restates version " + std::to_string(kMajor));
        uint16_t count = rd<uint16_t>(h + 12);
        pos += kHeader;
        std::vector<GridEntry> seg(count);
        for (uint16_t i = 0; i < count; ++i) {
            unsigned char m[kMeta];
            std::fseek(f, long(pos), SEEK_SET);
            if (std::fread(m, 1, kMeta, f) != kMeta) return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": Failed reading GridMetaData");
            GridEntry &g = seg[i];
            g.grid_size = rd<uint64_t>(m + kMdGridSize);
            g.file_size = rd<uint64_t>(m + kMdFileSize);
            g.grid_type = rd<uint32_t>(m + kMdGridType);
            g.grid_class = rd<uint32_t>(m + kMdGridClass);
            g.codec = rd<uint16_t>(m + kMdCodec);
            uint32_t name_size = rd<uint32_t>(m + kMdNameSize);
            if (name_size > 4096 || pos + kMeta + name_size > uint64_t(file_bytes))
                return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": Failed reading GridMetaData");
            std::vector<char> nm(name_size + 1, 0);
            if (name_size && std::fread(nm.data(), 1, name_size, f) != name_size)
                return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": Failed reading GridMetaData");
            g.name = nm.data();
            pos += kMeta + name_size;
        }
        for (GridEntry &g : seg) {
            g.data_pos = pos;
            if (g.file_size > uint64_t(file_bytes) - pos) return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": truncated grid data");
            pos += g.file_size;
            out->push_back(g);
        }
    }
    return AVR_OK;
}

avr_status load_buffer(FILE *f, const std::string &path, const GridEntry &g, std::vector<unsigned char> *buf) {
    if (g.grid_size < kGridData + kTreeData + kRootData || g.grid_size > (uint64_t(1) << 40))
        return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": implausible grid size");
    buf->resize(g.grid_size);
    std::fseek(f, long(g.data_pos), SEEK_SET);
    if (g.codec == 0) {
        if (g.file_size != g.grid_size || std::fread(buf->data(), 1, g.grid_size, f) != g.grid_size)
            return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": Failed to read Tree from file");
        return AVR_OK;
    }
    if (g.codec == 1) {  // ZIP: u64 compressed size + one zlib stream (nanovdb::io::Internal::write)
        unsigned char sz[8];
        if (g.file_size < 8 || std::fread(sz, 1, 8, f) != 8) return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": truncated ZIP grid");
        uint64_t n = rd<uint64_t>(sz);
        if (n + 8 > g.file_size) return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": truncated ZIP grid");
        std::vector<unsigned char> z(n);
        if (std::fread(z.data(), 1, n, f) != n) return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": truncated ZIP grid");
        void *lib = dlopen("libz.so.1", RTLD_NOW);
        using uncompress_t = int (*)(unsigned char *, unsigned long *, const unsigned char *, unsigned long);
        uncompress_t un = lib ? (uncompress_t)dlsym(lib, "uncompress") : nullptr;
        if (!un) return fail(AVR_ERR_UNSUPPORTED, "nanovdb: " + path + ": ZIP codec needs libz.so.1");
        unsigned long got = (unsigned long)g.grid_size;
        if (un(buf->data(), &got, z.data(), (unsigned long)n) != 0 || got != g.grid_size)
            return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": Internal read error in ZIP");
        return AVR_OK;
    }
    return fail(AVR_ERR_UNSUPPORTED, "nanovdb: " + path + ": BLOSC compression codec is not supported by this reader");
}

struct Walker {
    const unsigned char *base;
    size_t size;
    int32_t lo[3], hi[3];  // inclusive index box
    int64_t dims[3];
    float *out;
    bool ok = true;
    // sparse sink (avr_nvdb_read_blocks): 8^3 blocks instead of the dense array
    bool sparse = false, too_many = false;
    std::vector<int32_t> *origins = nullptr;
    std::vector<float> *blocks = nullptr;
    size_t tile_blocks = 0;
    static constexpr size_t kMaxTileBlocks = size_t(1) << 22;

    bool in_buf(const unsigned char *p, size_t n) const { return p >= base && size_t(p - base) + n <= size; }
    // constant value over [o, o + side)^3 clipped to the box
    void fill(const int32_t o[3], int32_t side, float v) {
        if (v == 0.f) return;  // the dense array starts at the background (0)
        int32_t a[3], b[3];
        for (int i = 0; i < 3; ++i) {
            a[i] = o[i] > lo[i] ? o[i] : lo[i];
            int64_t e = int64_t(o[i]) + side - 1;
            b[i] = e < hi[i] ? int32_t(e) : hi[i];
            if (a[i] > b[i]) return;
        }
        if (sparse) {
            // the tile cut into the 8^3 blocks that meet the index box (tiles are aligned to their own extent >= 8)
            for (int32_t z = a[2] & ~7; z <= b[2] && !too_many; z += 8)
                for (int32_t y = a[1] & ~7; y <= b[1] && !too_many; y += 8)
                    for (int32_t x = a[0] & ~7; x <= b[0]; x += 8) {
                        if (++tile_blocks > kMaxTileBlocks) { too_many = true; break; }
                        origins->insert(origins->end(), {x, y, z});
                        blocks->insert(blocks->end(), 512, v);
                    }
            return;
        }
        for (int32_t z = a[2]; z <= b[2]; ++z)
            for (int32_t y = a[1]; y <= b[1]; ++y) {
                float *row = out + ((int64_t(z) - lo[2]) * dims[1] + (y - lo[1])) * dims[0];
                for (int32_t x = a[0]; x <= b[0]; ++x) row[x - lo[0]] = v;
            }
    }
    void leaf(const unsigned char *p, const int32_t o[3]) {
        if (!in_buf(p, kLeafSize)) { ok = false; return; }
        const unsigned char *vals = p + kLeafValues;
        if (sparse) {
            // the leaf as it is, transposed to x-fastest; leaves that hold nothing but zeros are dropped
            float tmp[512];
            bool any = false;
            for (int x = 0; x < 8; ++x)
                for (int y = 0; y < 8; ++y)
                    for (int z = 0; z < 8; ++z) {
                        float v = rd<float>(vals + 4 * ((x << 6) | (y << 3) | z));
                        tmp[x + 8 * y + 64 * z] = v;
                        any |= v != 0.f;
                    }
            if (any) {
                origins->insert(origins->end(), {o[0], o[1], o[2]});
                blocks->insert(blocks->end(), tmp, tmp + 512);
            }
            return;
        }
        for (int x = 0; x < 8; ++x) {
            int32_t X = o[0] + x;
            if (X < lo[0] || X > hi[0]) continue;
            for (int y = 0; y < 8; ++y) {
                int32_t Y = o[1] + y;
                if (Y < lo[1] || Y > hi[1]) continue;
                for (int z = 0; z < 8; ++z) {
                    int32_t Z = o[2] + z;
                    if (Z < lo[2] || Z > hi[2]) continue;
                    out[((int64_t(Z) - lo[2]) * dims[1] + (Y - lo[1])) * dims[0] + (X - lo[0])] = rd<float>(vals + 4 * ((x << 6) | (y << 3) | z));
                }
            }
        }
    }
    // internal node: log2dim 4 (lower, children = leaves, child extent 8) or 5 (upper, children = lower nodes, 128)
    void internal(const unsigned char *p, const int32_t o[3], int log2dim) {
        const bool upper = log2dim == 5;
        const size_t node_size = upper ? kUpperSize : kLowerSize, mask_off = upper ? kUpperChildMask : kLowerChildMask,
                     table_off = upper ? kUpperTable : kLowerTable;
        if (!in_buf(p, node_size)) { ok = false; return; }
        const int dim = 1 << log2dim, child_extent = upper ? 128 : 8;
        for (int x = 0; x < dim; ++x)
            for (int y = 0; y < dim; ++y)
                for (int z = 0; z < dim; ++z) {
                    int32_t co[3] = {o[0] + x * child_extent, o[1] + y * child_extent, o[2] + z * child_extent};
                    // outside the box: nothing to write
                    if (co[0] > hi[0] || co[1] > hi[1] || co[2] > hi[2] || int64_t(co[0]) + child_extent - 1 < lo[0] ||
                        int64_t(co[1]) + child_extent - 1 < lo[1] || int64_t(co[2]) + child_extent - 1 < lo[2])
                        continue;
                    uint32_t n = (uint32_t(x) << (2 * log2dim)) | (uint32_t(y) << log2dim) | uint32_t(z);
                    uint64_t word = rd<uint64_t>(p + mask_off + 8 * (n >> 6));
                    const unsigned char *entry = p + table_off + 8 * size_t(n);
                    if (word >> (n & 63) & 1) {
                        int64_t off = rd<int64_t>(entry);
                        if (upper) internal(p + off, co, 4); else leaf(p + off, co);
                        if (!ok) return;
                    } else
                        fill(co, child_extent, rd<float>(entry));
                }
    }
};

avr_status read_grid(const std::string &path, const char *grid_name, avr_nanovdb_grid *grid, avr_nvdb_info *info, bool sparse) {
    FILE *f = std::fopen(path.c_str(), "rb");
    if (!f) return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": Unable to open file named \"" + path + "\" for input");
    std::vector<GridEntry> grids;
    avr_status st = list_grids(f, path, &grids);
    if (st != AVR_OK) { std::fclose(f); return st; }
    const GridEntry *pick = nullptr;
    for (const GridEntry &g : grids)
        if (!grid_name || !*grid_name || g.name == grid_name) { pick = &g; break; }
    if (info) {
        std::memset(info, 0, sizeof(*info));
        info->grids_in_file = int32_t(grids.size());
    }
    if (!pick) {
        std::fclose(f);
        std::memset(grid, 0, sizeof(*grid));   // values == NULL: "grid absent", as an empty GridHandle
        return AVR_OK;
    }
    if (pick->grid_type != 1) { std::fclose(f); return fail(AVR_ERR_UNSUPPORTED, "nanovdb: " + path + ": \"" + pick->name + "\" is not a FloatGrid"); }
    // GridClass: 0 Unknown, 1 LevelSet, 2 FogVolume, ... (media.cpp:501-502 accepts FogVolume and Unknown)
    if (pick->grid_class != 2 && pick->grid_class != 0) {
        std::fclose(f);
        return fail(AVR_ERR_BAD_ARG, path + ": \"" + pick->name + "\" isn't a FogVolume grid?");
    }
    std::vector<unsigned char> buf;
    st = load_buffer(f, path, *pick, &buf);
    std::fclose(f);
    if (st != AVR_OK) return st;
    const unsigned char *b = buf.data();
    if (rd<uint64_t>(b) != kMagic || int(rd<uint32_t>(b + kGdVersion) >> 21) != kMajor || rd<uint32_t>(b + kGdGridType) != 1)
        return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": grid buffer is not a version-" + std::to_string(kMajor) + " FloatGrid");
    const unsigned char *tree = b + kGridData;
    uint64_t off_root = rd<uint64_t>(tree + 24);
    const unsigned char *root = tree + off_root;
    if (size_t(root - b) + kRootData > buf.size()) return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": corrupt tree offsets");
    Walker w;
    w.base = b; w.size = buf.size();
    for (int i = 0; i < 3; ++i) { w.lo[i] = rd<int32_t>(root + 4 * i); w.hi[i] = rd<int32_t>(root + 12 + 4 * i); }
    uint32_t table_size = rd<uint32_t>(root + 24);
    float background = rd<float>(root + 28);
    if (background != 0.f) return fail(AVR_ERR_UNSUPPORTED, "nanovdb: " + path + ": background value must be 0 (fog volume)");
    int64_t total = 1;
    for (int i = 0; i < 3; ++i) {
        w.dims[i] = int64_t(w.hi[i]) - w.lo[i] + 1;
        if (w.dims[i] < 1 || w.dims[i] > (1 << 20)) return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": empty or implausible index bounding box");
        total *= w.dims[i];
    }
    if (!sparse && total > (int64_t(1) << 36)) return fail(AVR_ERR_OOM, "nanovdb: " + path + ": index bounding box too large to flatten");
    float *values = nullptr;
    std::vector<int32_t> origins;
    std::vector<float> blocks;
    if (sparse) {
        w.sparse = true; w.origins = &origins; w.blocks = &blocks;
    } else {
        values = (float *)std::calloc(size_t(total), sizeof(float));
        if (!values) return fail(AVR_ERR_OOM, "nanovdb: out of memory flattening the grid");
    }
    w.out = values;
    if (size_t(root - b) + kRootData + size_t(table_size) * kRootTile > buf.size()) { std::free(values); return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": corrupt root table"); }
    for (uint32_t t = 0; t < table_size && w.ok && !w.too_many; ++t) {
        const unsigned char *tile = root + kRootData + size_t(t) * kRootTile;
        uint64_t key = rd<uint64_t>(tile);
        int64_t child = rd<int64_t>(tile + 8);
        // KeyToCoord: x is the upper 21 bits, z the lower; << 12 (the upper node's extent)
        int32_t o[3] = {int32_t(uint32_t((key >> 42) & 0x1fffff) << 12), int32_t(uint32_t((key >> 21) & 0x1fffff) << 12),
                        int32_t(uint32_t(key & 0x1fffff) << 12)};
        if (child != 0) w.internal(root + child, o, 5);
        else w.fill(o, 4096, rd<float>(tile + 20));
    }
    if (!w.ok) { std::free(values); return fail(AVR_ERR_BAD_ARG, "nanovdb: " + path + ": node offset outside the grid buffer"); }
    if (w.too_many)
        return fail(AVR_ERR_UNSUPPORTED, "nanovdb: " + path + ": value tiles expand to more than 2^22 blocks (read the grid flattened instead)");
    std::memset(grid, 0, sizeof(*grid));
    grid->values = values;
    if (sparse) {
        const size_t nb = origins.size() / 3;
        int32_t *po = (int32_t *)std::malloc(std::max<size_t>(1, nb) * 3 * sizeof(int32_t));
        float *pv = (float *)std::malloc(std::max<size_t>(1, nb) * 512 * sizeof(float));
        if (!po || !pv) { std::free(po); std::free(pv); return fail(AVR_ERR_OOM, "nanovdb: out of memory collecting the leaves"); }
        if (nb) {
            std::memcpy(po, origins.data(), nb * 3 * sizeof(int32_t));
            std::memcpy(pv, blocks.data(), nb * 512 * sizeof(float));
        }
        grid->n_blocks = int64_t(nb);
        grid->block_origin = po;
        grid->block_values = pv;
    }
    for (int i = 0; i < 3; ++i) {
        grid->index_min[i] = w.lo[i];
        grid->dims[i] = int32_t(w.dims[i]);
        grid->vec[i] = rd<float>(b + kGdMap + kMapVecF + 4 * i);
    }
    for (int i = 0; i < 9; ++i) grid->inv_mat[i] = rd<float>(b + kGdMap + kMapInvMatF + 4 * i);
    for (int i = 0; i < 6; ++i) grid->world_bbox[i] = rd<double>(b + kGdWorldBBox + 8 * i);
    if (info) {
        info->leaf_nodes = rd<uint32_t>(tree + 32); info->lower_nodes = rd<uint32_t>(tree + 36); info->upper_nodes = rd<uint32_t>(tree + 40);
        info->active_voxels = int64_t(rd<uint64_t>(tree + 56));
        info->grid_bytes = int64_t(pick->grid_size);
        info->codec = pick->codec;
        std::strncpy(info->grid_name, pick->name.c_str(), sizeof(info->grid_name) - 1);
    }
    return AVR_OK;
}

}  // namespace

extern "C" avr_status avr_nvdb_read_grid(const char *path, const char *grid_name, avr_nanovdb_grid *grid, avr_nvdb_info *info) {
    if (!path || !grid) return fail(AVR_ERR_BAD_ARG, "avr_nvdb_read_grid: NULL argument");
    try {
        return read_grid(path, grid_name, grid, info, false);
    } catch (const std::bad_alloc &) {
        return fail(AVR_ERR_OOM, "avr_nvdb_read_grid: out of memory");
    }
}

extern "C" avr_status avr_nvdb_read_blocks(const char *path, const char *grid_name, avr_nanovdb_grid *grid, avr_nvdb_info *info) {
    if (!path || !grid) return fail(AVR_ERR_BAD_ARG, "avr_nvdb_read_blocks: NULL argument");
    try {
        return read_grid(path, grid_name, grid, info, true);
    } catch (const std::bad_alloc &) {
        return fail(AVR_ERR_OOM, "avr_nvdb_read_blocks: out of memory");
    }
}

extern "C" void avr_nvdb_free_grid(avr_nanovdb_grid *grid) {
    if (!grid) return;
    std::free((void *)grid->values);
    std::free((void *)grid->block_origin);
    std::free((void *)grid->block_values);
    grid->values = nullptr;
    grid->block_origin = nullptr;
    grid->block_values = nullptr;
    grid->n_blocks = 0;
}

extern "C" avr_status avr_medium_create_nanovdb_file(avr_ctx *ctx, const char *filename, const char *gridname,
                                                     const char *temperaturename, const avr_nanovdb_desc *params, int layout,
                                                     avr_medium **out) {
    if (!ctx || !params || !out) return avr::set_err(ctx, AVR_ERR_BAD_ARG, "avr_medium_create_nanovdb_file: NULL argument");
    *out = nullptr;
    // NanoVDBMedium::Create, media.cpp:631-633
    if (!filename || !*filename) return avr::set_err(ctx, AVR_ERR_BAD_ARG, "Must supply \"filename\" to \"nanovdb\" medium.");
    avr_nanovdb_desc d = *params;
    std::memset(&d.density, 0, sizeof(d.density));
    std::memset(&d.temperature, 0, sizeof(d.temperature));
    avr_status st = avr_nvdb_read_grid(filename, temperaturename && *temperaturename ? temperaturename : "temperature", &d.temperature, nullptr);
    if (st != AVR_OK) return avr::set_err(ctx, st, avr_last_error(nullptr));
    // the brick pool is built from the leaves when nothing else needs the box (AVR_NVDB_FLATTEN=1: the flattened
    // route, for comparisons)
    const char *env = std::getenv("AVR_NVDB_FLATTEN");
    bool from_leaves = layout == AVR_LAYOUT_BRICK_F16 && !d.temperature.values && !(env && *env == '1');
    const char *dname = gridname && *gridname ? gridname : "density";
    if (from_leaves) {
        st = avr_nvdb_read_blocks(filename, dname, &d.density, nullptr);
        if (st == AVR_ERR_UNSUPPORTED) from_leaves = false;   // huge value tiles: flatten after all
        else if (st != AVR_OK) { avr_nvdb_free_grid(&d.temperature); return avr::set_err(ctx, st, avr_last_error(nullptr)); }
    }
    if (!from_leaves) {
        st = avr_nvdb_read_grid(filename, dname, &d.density, nullptr);
        if (st != AVR_OK) { avr_nvdb_free_grid(&d.temperature); return avr::set_err(ctx, st, avr_last_error(nullptr)); }
    }
    if (!d.density.values && !d.density.block_values) {
        avr_nvdb_free_grid(&d.temperature);
        return avr::set_err(ctx, AVR_ERR_BAD_ARG, std::string(filename) + ": didn't find \"density\" grid.");
    }
    st = avr_medium_create_nanovdb(ctx, &d, layout, out);
    avr_nvdb_free_grid(&d.density);
    avr_nvdb_free_grid(&d.temperature);
    return st;
}
