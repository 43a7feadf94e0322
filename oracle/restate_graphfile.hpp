// oracle/restate_graphfile.hpp — TEST INFRASTRUCTURE ONLY (see restate.hpp header).
// src/graph containers and text format: Graph::AddVertex / AddEdge (graph.cpp:150-235), WriteToStream /
// ReadFromStream (:284-416), FreeGraph stream header (:644-655), and
// LightingCalculator::GetTransportMatrix (lighting_calculator.cpp:61-82).
// Records keep their insertion order (the reference iterates unordered_maps, whose order is an
// implementation detail; callers that compare files feed the order the reference's file has).
#pragma once
#include <fstream>
#include <map>
#include <string>
#include <unordered_map>

#include "restate_graph.hpp"

namespace avro {

struct GVertex {
    int id;
    float p[3];
    int type = 0;  // RayVertexType none
    float light = -1;
    int samples = 0;
    float range = -1;
    bool has_coors = false;
    int coors[3] = {0, 0, 0};
    std::map<int, int> outEdges, inEdges;  // other vertex id -> index into edges
};
struct GEdge {
    int id, from, to, samples;
};
struct GPath {
    int id;
    std::vector<int> verts;
};

struct FreeGraphR {
    float vertex_radius = -1;
    bool flags[5] = {false, false, false, false, false};  // coors, samples, types, lighting, searchRange
    int curVertexId = 0, curEdgeId = 0, curPathId = 0;
    std::vector<GVertex> vertices;
    std::vector<GEdge> edges;
    std::vector<GPath> paths;
    std::unordered_map<int, size_t> vindex, eindex;

    // Graph::AddVertex(id, p, data): FreeGraph::AddVertex(id, p, data, incrId=false) (graph.cpp)
    GVertex &add_vertex(int id, const float p[3]) {
        vindex[id] = vertices.size();
        GVertex v;
        v.id = id;
        for (int i = 0; i < 3; ++i) v.p[i] = p[i];
        vertices.push_back(v);
        return vertices.back();
    }
    // Graph::AddEdge(id, from, to, data, incrId), graph.cpp:202-235
    void add_edge(int id, int from, int to, int samples) {
        auto f = vindex.find(from), t = vindex.find(to);
        if (f == vindex.end() || t == vindex.end()) return;
        GVertex &vf = vertices[f->second], &vt = vertices[t->second];
        auto er = vt.inEdges.find(from);
        if (er != vt.inEdges.end() && vf.outEdges.find(to) != vf.outEdges.end()) {
            edges[size_t(er->second)].samples += samples;  // EdgeData::MergeWithDataFrom
            return;
        }
        int idx = int(edges.size());
        edges.push_back({id, from, to, samples});
        eindex[id] = size_t(idx);
        vf.outEdges.emplace(to, idx);
        vt.inEdges.emplace(from, idx);
    }

    // Graph::WriteToDisk + FreeGraph::WriteToStream + Graph::WriteToStream
    void write(const std::string &path, const std::string &desc, const bool opts[3]) const {
        std::ofstream out(path);
        const char SEP = ' ', NEW = '\n';
        out << desc << NEW;
        out << "free" << SEP << vertex_radius << NEW;
        for (int i = 0; i < 5; ++i) out << (flags[i] ? "True" : "False") << (i < 4 ? SEP : NEW);
        size_t nv = opts[0] ? vertices.size() : 0, ne = opts[1] ? edges.size() : 0, np = opts[2] ? paths.size() : 0;
        out << curVertexId << SEP << curEdgeId << SEP << curPathId << SEP << nv << SEP << ne << SEP << np << NEW;
        if (opts[0])
            for (const GVertex &v : vertices) {
                out << v.id << SEP << v.p[0] << SEP << v.p[1] << SEP << v.p[2] << SEP;
                if (flags[2]) out << v.type << SEP;
                if (flags[3]) out << v.light << SEP;
                if (flags[1]) out << v.samples << SEP;
                if (flags[4]) out << v.range << SEP;
                if (flags[0]) out << v.coors[0] << SEP << v.coors[1] << SEP << v.coors[2] << SEP;
                out << NEW;
            }
        if (opts[1])
            for (const GEdge &e : edges) {
                out << e.id << SEP << e.from << SEP << e.to << SEP;
                if (flags[1]) out << e.samples << SEP;
                out << NEW;
            }
        if (opts[2])
            for (const GPath &p : paths) {
                out << p.id << SEP << p.verts.size() << SEP;
                for (int v : p.verts) out << v << SEP;
                out << NEW;
            }
    }

    // FreeGraph::ReadFromStream + Graph::ReadFromStream
    bool read(const std::string &path) {
        std::ifstream in(path);
        if (!in) return false;
        std::string description, name, f[5];
        in >> description >> name >> vertex_radius;
        in >> f[0] >> f[1] >> f[2] >> f[3] >> f[4];
        for (int i = 0; i < 5; ++i) flags[i] = f[i] == "True";
        int nv, ne, np;
        in >> curVertexId >> curEdgeId >> curPathId >> nv >> ne >> np;
        for (int i = 0; i < nv; ++i) {
            int id;
            float p[3];
            in >> id >> p[0] >> p[1] >> p[2];
            int type = -1, samples = -1;
            float light = -1, range = -1;
            if (flags[2]) in >> type;
            if (flags[3]) in >> light;
            if (flags[1]) in >> samples;
            if (flags[4]) in >> range;
            GVertex &v = add_vertex(id, p);
            v.type = type; v.light = light; v.samples = samples; v.range = range;
            if (flags[0]) { in >> v.coors[0] >> v.coors[1] >> v.coors[2]; v.has_coors = true; }
        }
        for (int i = 0; i < ne; ++i) {
            int id, from, to, samples = -1;
            in >> id >> from >> to;
            if (flags[1]) in >> samples;
            add_edge(id, from, to, samples);
        }
        for (int i = 0; i < np; ++i) {
            GPath p;
            int len;
            in >> p.id >> len;
            for (int j = 0; j < len; ++j) { int v; in >> v; p.verts.push_back(v); }
            paths.push_back(p);
        }
        return bool(in);
    }

    // GetTransportMatrix: rows = vertex ids 0..n-1 (CheckSequentialIds), per row the std::map outEdges
    // in key order: (col = other vertex id, val = float(edge samples) / float(vertex samples))
    Csr transport() const {
        Csr T;
        T.n = int(vertices.size());
        T.rowptr.assign(size_t(T.n) + 1, 0);
        std::vector<const GVertex *> by_id(size_t(T.n), nullptr);
        for (const GVertex &v : vertices) by_id[size_t(v.id)] = &v;
        for (int r = 0; r < T.n; ++r) {
            const GVertex &v = *by_id[size_t(r)];
            for (auto &oe : v.outEdges) {
                T.col.push_back(oe.first);
                T.val.push_back(float(edges[size_t(oe.second)].samples) / float(v.samples));
            }
            T.rowptr[size_t(r) + 1] = int64_t(T.col.size());
        }
        return T;
    }
};

}  // namespace avro
