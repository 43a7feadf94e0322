"""CPU (host-only entry points of the C ABI, no GPU needed): the graph text-file reader / writer of
libavr_b200.so against the reference's own Graph::WriteToStream / ReadFromStream output.

The golden fixture (tests/golden, case "graphfile") holds, for files written by the fork's FreeGraph
compiled from /root/reference: the 4 header lines as bytes, the record lines as a sorted byte blob (the
reference walks unordered_maps, so record order is an implementation detail) and what its reader
parses back.  Bar: byte-identical files, bit-identical parsed arrays.
"""
import os

import numpy as np
import pytest

from acceleratedvolrenderer_b200 import scenes
from acceleratedvolrenderer_b200.api import AvrError, GraphFile

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.npz")
VARIANTS = (("maker", (0, 0, 0, 1, 1), (1, 0, 0)), ("full", (1, 1, 1, 1, 1), (1, 1, 1)), ("bare", (0, 0, 0, 0, 0), (1, 1, 0)))


def build(orc, n, ne):
    """The golden case's graph in the oracle, after GetTransportMatrix + ComputeFinalLight (which stores
    lightScalar into the vertices, lighting_calculator.cpp:55-56)."""
    r = scenes.random_edge_graph(n, ne, seed=n)
    g = orc.graph_build(0.0125, np.arange(n), r["pos"], vtype=r["vtype"], vlight=r["vlight"], vsamples=r["vsamples"],
                        vrange=r["vrange"], vcoors=r["vcoors"], eid=np.arange(r["efrom"].size), efrom=r["efrom"],
                        eto=r["eto"], esamples=r["esamples"], pid=np.arange(r["poff"].size - 1), poff=r["poff"],
                        pverts=r["pverts"])
    rng = np.random.default_rng(12)
    for _n, _ne in ((40, 150), (1500, 7000)):  # same L0 stream as golden_cases.case_graphfile
        L0 = rng.random(_n, dtype=np.float32)
        if _n == n:
            break
    for b in (0, 3):
        g.transport(L0, b)
    return g, r


def split(data):
    lines = data.split(b"\n")
    return b"\n".join(lines[:4]), b"\n".join(sorted(lines[4:]))


@pytest.mark.parametrize("n,ne", [(40, 150), (1500, 7000)])
def test_writer_bytes_match_reference(orc, tmp_path, n, ne):
    golden = np.load(GOLDEN)
    g, r = build(orc, n, ne)
    st = g.get()  # merged edges, updated lightScalar: the state the reference wrote
    g.close()
    gf = GraphFile(0.0125, st["vid"], st["pos"], vertex_type=st["vtype"], vertex_light=st["vlight"],
                   vertex_samples=st["vsamples"], vertex_range=st["vrange"], vertex_coors=st["vcoors"],
                   edge_id=st["eid"], edge_from=st["efrom"], edge_to=st["eto"], edge_samples=st["esamples"],
                   path_id=np.arange(r["poff"].size - 1), path_offset=r["poff"], path_vertices=r["pverts"])
    for tag, flags, opts in VARIANTS:
        path = tmp_path / f"{tag}.txt"
        gf.write(path, flags, opts)
        header, records = split(path.read_bytes())
        assert header == golden[f"graphfile/g{n}_{tag}_header"].tobytes(), tag
        assert records == golden[f"graphfile/g{n}_{tag}_records"].tobytes(), tag


@pytest.mark.parametrize("n,ne", [(40, 150), (1500, 7000)])
def test_reader_matches_reference(orc, tmp_path, n, ne):
    golden = np.load(GOLDEN)
    g, r = build(orc, n, ne)
    for tag, flags, opts in VARIANTS:
        path = tmp_path / f"{tag}.txt"
        g.write(str(path), "basic", flags, opts)  # oracle writer, pinned byte-exact to the reference's
        gf = GraphFile.read(path)
        want = {k.split("_read_")[1]: golden[k] for k in golden.files if k.startswith(f"graphfile/g{n}_{tag}_read_")}
        counts = golden[f"graphfile/g{n}_{tag}_counts"]
        assert tuple(gf.flags) == tuple(counts[:5]) and gf.cur_ids == tuple(counts[5:8])
        assert gf.counts[0] == counts[8] and np.float32(gf.vertex_radius) == golden[f"graphfile/g{n}_{tag}_radius"][0]
        order = np.argsort(gf.vertex_id, kind="stable")
        assert np.array_equal(gf.vertex_id[order], want["vid"])
        assert np.array_equal(gf.vertex_pos[order].view(np.uint32), want["pos"].view(np.uint32))
        if flags[3]:
            assert np.array_equal(gf.vertex_light[order].view(np.uint32), want["vlight"].view(np.uint32))
        else:
            assert gf.vertex_light is None
        if flags[4]:
            assert np.array_equal(gf.vertex_range[order].view(np.uint32), want["vrange"].view(np.uint32))
        if flags[1]:
            assert np.array_equal(gf.vertex_samples[order], want["vsamples"])
        if flags[2]:
            assert np.array_equal(gf.vertex_type[order], want["vtype"])
        if flags[0]:
            assert np.array_equal(gf.vertex_coors[order], want["vcoors"])
        if opts[1]:
            eo = np.argsort(gf.edge_id, kind="stable")
            assert np.array_equal(gf.edge_id[eo], want["eid"]) and np.array_equal(gf.edge_from[eo], want["efrom"])
            assert np.array_equal(gf.edge_to[eo], want["eto"])
            if flags[1]:
                assert np.array_equal(gf.edge_samples[eo], want["esamples"])
        else:
            assert gf.counts[1] == 0
        if opts[2]:
            assert gf.counts[2] == r["poff"].size - 1
            assert np.array_equal(gf.path_offset, r["poff"]) and np.array_equal(gf.path_vertices, r["pverts"])
    g.close()


def test_round_trip_and_errors(tmp_path):
    rng = np.random.default_rng(0)
    n = 1000
    gf = GraphFile(0.25, np.arange(n), rng.normal(size=(n, 3)) * 1e3, vertex_light=rng.random(n) * 1e-7,
                   vertex_range=rng.random(n), description="grid_lighting", cur_ids=(n, 7, 9))
    p1, p2 = tmp_path / "a.txt", tmp_path / "b.txt"
    gf.write(p1, (0, 0, 0, 1, 1), (1, 0, 0))
    back = GraphFile.read(p1)
    assert back.description == "grid_lighting" and back.cur_ids == (n, 7, 9)
    # 6 significant digits survive; a second write of the parsed values is byte-identical (idempotence)
    assert np.allclose(back.vertex_pos, gf.vertex_pos, rtol=1e-5)
    back.write(p2, (0, 0, 0, 1, 1), (1, 0, 0))
    assert p1.read_bytes() == p2.read_bytes()
    empty = GraphFile(0.5, np.zeros(0, np.int32), np.zeros((0, 3), np.float32))
    empty.write(tmp_path / "e.txt", (0, 0, 0, 0, 0), (1, 1, 1))
    e = GraphFile.read(tmp_path / "e.txt")
    assert e.counts == (0, 0, 0) and e.vertex_id is None
    with pytest.raises(AvrError, match="cannot open"):
        GraphFile.read(tmp_path / "missing.txt")
    (tmp_path / "trunc.txt").write_text("basic\nfree 0.5\nFalse False False False False\n0 0 0 3 0 0\n0 1 2 3 \n1 1 2")
    with pytest.raises(AvrError, match="truncated"):
        GraphFile.read(tmp_path / "trunc.txt")
    (tmp_path / "uni.txt").write_text("basic\nuniform 0.5\n")
    with pytest.raises(AvrError, match="free"):
        GraphFile.read(tmp_path / "uni.txt")
    with pytest.raises(AvrError, match="flag is set"):
        gf.write(tmp_path / "x.txt", (0, 1, 0, 1, 1), (1, 0, 0))  # useSamples without vertex samples


# ---- binary ".avrg" sidecar (include/avr.h: avr_graph_sidecar) -------------------------
def sidecar_case(orc, n=1500, ne=7000):
    from acceleratedvolrenderer_b200.api import GraphSidecar
    g, r = build(orc, n, ne)
    T = g.transport()  # the reference-pinned GetTransportMatrix of the same graph (rowptr, col, val)
    g.close()
    rng = np.random.default_rng(5)
    return GraphSidecar(T[0], T[1], T[2], vertex_pos=r["pos"], light=rng.random(n, dtype=np.float32),
                        search_range=r["vrange"]), T


def test_reader_survives_hostile_headers(tmp_path):
    """Header counts a file of that size cannot hold, overflowing integers: an error status, never an allocation of
    gigabytes or an exception through the C boundary (the library never aborts the host)."""
    for counts in ("0 0 0 2000000000 0 0", "0 0 0 3 2000000000 0", "0 0 0 3 0 2000000000", "0 0 0 99999999999999999999 0 0"):
        p = tmp_path / "bad.txt"
        p.write_text(f"basic\nfree 0.1\nFalse False False False False\n{counts}\n0 0 0 0\n1 1 1 1\n2 2 2 2\n")
        with pytest.raises(AvrError):
            GraphFile.read(p)
    ok = tmp_path / "ok.txt"
    ok.write_text("basic\nfree 0.1\nFalse False False False False\n3 0 0 3 0 0\n0 0 0 0\n1 1 1 1\n2 2 2 2\n")
    assert GraphFile.read(ok).vertex_pos.shape == (3, 3)


def test_sidecar_round_trip_is_bit_exact(orc, tmp_path):
    from acceleratedvolrenderer_b200.api import GraphSidecar
    sc, T = sidecar_case(orc)
    p = tmp_path / "graph.avrg"
    sc.write(p)
    n, nnz = sc.n_vertices, sc.col.size
    assert os.path.getsize(p) == 24 + 8 * (n + 1) + 8 * nnz + 20 * n + 8
    back = GraphSidecar.read(p)
    for name in ("rowptr", "col", "val", "vertex_pos", "light", "search_range"):
        a, b = getattr(sc, name), getattr(back, name)
        assert a.reshape(-1).tobytes() == b.reshape(-1).tobytes(), name
    # optional arrays may be absent; an edgeless graph is legal
    bare = GraphSidecar(np.zeros(4, np.int64), np.zeros(0, np.int32), np.zeros(0, np.float32))
    bare.write(p)
    b2 = GraphSidecar.read(p)
    assert b2.n_vertices == 3 and b2.col.size == 0 and b2.vertex_pos is None and b2.light is None and b2.search_range is None


def test_sidecar_rejects_damaged_files(orc, tmp_path):
    from acceleratedvolrenderer_b200.api import GraphSidecar
    sc, _ = sidecar_case(orc, 40, 150)
    p = tmp_path / "graph.avrg"
    sc.write(p)
    data = bytearray(open(p, "rb").read())
    cases = {"truncated": bytes(data[:-9]), "magic": b"AVRX" + bytes(data[4:]), "version": bytes(data[:4]) + b"\x02\0\0\0" + bytes(data[8:])}
    flipped = bytearray(data); flipped[len(data) // 2] ^= 0x10
    cases["bit flip"] = bytes(flipped)
    for name, blob in cases.items():
        q = tmp_path / "bad.avrg"
        open(q, "wb").write(blob)
        with pytest.raises(AvrError):
            GraphSidecar.read(q)
    with pytest.raises(AvrError):
        GraphSidecar.read(tmp_path / "missing.avrg")
    # a CSR that does not describe n vertices is refused at write time
    bad = GraphSidecar(np.array([0, 2, 1], np.int64), np.zeros(1, np.int32), np.ones(1, np.float32))
    with pytest.raises(AvrError):
        bad.write(tmp_path / "bad2.avrg")


@pytest.mark.gpu
def test_sidecar_feeds_the_propagation(ctx, orc, tmp_path):
    """graph_maker -> .avrg -> kernel 3: the light the device propagates from a sidecar read back from disk
    is bit-identical to the oracle's ComputeFinalLight on the in-memory graph."""
    from acceleratedvolrenderer_b200.api import GraphSidecar
    from cpu_checker import Checker  # noqa: F401
    sc, T = sidecar_case(orc)
    p = tmp_path / "graph.avrg"
    sc.write(p)
    back = GraphSidecar.read(p)
    g = ctx.create_graph(back.rowptr, back.col, back.val)
    total, it = g.compute_final_light(back.light, 6)
    g.close()
    want, wit = orc.graph_propagate(T[0], T[1], T[2], sc.light, 6)
    assert it == wit and total.tobytes() == want.tobytes()
