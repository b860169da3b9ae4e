"""Gmsh MSH 4.1 reader (arcanefem_b200/msh.py), the stand-in for Arcane's mesh reader on the reference's case meshes
(SURVEY.md §8f-1).  A small mesh with physical groups is written here in both flavours of the published 4.1 layout (ASCII and
binary) and read back; when the reference tree is present (this container, not the GPU box) the reader is also pinned on
meshes/msh/circle_cut.msh through the committed fixture that tests/golden/make_golden.py made from it."""
import os
import struct

import numpy as np
import pytest

from arcanefem_b200 import mesh as gen
from arcanefem_b200 import msh

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def write_msh41(path, binary, node_tags, coords, blocks, physical_names, entities):
    """blocks: list of (entity_dim, entity_tag, element_type, int64 [n, 1 + nodes] records: element tag + node TAGS);
    entities: {dim: [(tag, [physical tags])]}; physical_names: [(dim, tag, name)]."""
    with open(path, "wb") as f:
        w = f.write
        w(f"$MeshFormat\n4.1 {1 if binary else 0} 8\n".encode())
        if binary:
            w(struct.pack("<i", 1) + b"\n")
        w(b"$EndMeshFormat\n$PhysicalNames\n")
        w(f"{len(physical_names)}\n".encode())
        for dim, tag, name in physical_names:
            w(f'{dim} {tag} "{name}"\n'.encode())
        w(b"$EndPhysicalNames\n$Entities\n")
        counts = [len(entities.get(d, [])) for d in range(4)]
        if binary:
            w(struct.pack("<4Q", *counts))
            for d in range(4):
                for tag, phys in entities.get(d, []):
                    w(struct.pack("<i", tag) + struct.pack(f"<{3 if d == 0 else 6}d", *([0.0] * (3 if d == 0 else 6))))
                    w(struct.pack("<Q", len(phys)) + struct.pack(f"<{len(phys)}i", *phys))
                    if d > 0:
                        w(struct.pack("<Q", 0))
            w(b"\n")
        else:
            w((" ".join(map(str, counts)) + "\n").encode())
            for d in range(4):
                for tag, phys in entities.get(d, []):
                    box = "0 0 0" if d == 0 else "0 0 0 1 1 1"
                    tail = "" if d == 0 else " 0"
                    w(f"{tag} {box} {len(phys)} {' '.join(map(str, phys))}{tail}\n".replace("  ", " ").encode())
        w(b"$EndEntities\n$Nodes\n")
        nn = len(node_tags)
        half = nn // 2  # two node blocks, to exercise the block loop
        parts = [(0, half), (half, nn)]
        if binary:
            w(struct.pack("<4Q", len(parts), nn, int(min(node_tags)), int(max(node_tags))))
            for a, b in parts:
                w(struct.pack("<3i", 2, 1, 0) + struct.pack("<Q", b - a))
                w(np.asarray(node_tags[a:b], dtype="<u8").tobytes())
                w(np.asarray(coords[a:b], dtype="<f8").tobytes())
            w(b"\n")
        else:
            w(f"{len(parts)} {nn} {min(node_tags)} {max(node_tags)}\n".encode())
            for a, b in parts:
                w(f"2 1 0 {b - a}\n".encode())
                for t in node_tags[a:b]:
                    w(f"{t}\n".encode())
                for x in coords[a:b]:
                    w(f"{float(x[0])!r} {float(x[1])!r} {float(x[2])!r}\n".encode())
        w(b"$EndNodes\n$Elements\n")
        ne = sum(b[3].shape[0] for b in blocks)
        if binary:
            w(struct.pack("<4Q", len(blocks), ne, 1, ne))
            for edim, etag, etype, rec in blocks:
                w(struct.pack("<3i", edim, etag, etype) + struct.pack("<Q", rec.shape[0]))
                w(np.asarray(rec, dtype="<u8").tobytes())
            w(b"\n")
        else:
            w(f"{len(blocks)} {ne} 1 {ne}\n".encode())
            for edim, etag, etype, rec in blocks:
                w(f"{edim} {etag} {etype} {rec.shape[0]}\n".encode())
                for r in rec:
                    w((" ".join(map(str, r)) + "\n").encode())
        w(b"$EndElements\n")


@pytest.mark.parametrize("binary", [False, True], ids=["ascii", "binary"])
def test_round_trip_of_a_small_mesh_with_groups(tmp_path, binary):
    cells, coords = gen.square(3)
    nn = coords.shape[0]
    rng = np.random.default_rng(4)
    tags = (rng.permutation(nn) * 3 + 7).astype(np.int64)  # sparse, unordered Gmsh tags: local id = rank of the tag
    order = rng.permutation(nn)  # nodes listed in yet another order in the file
    bottom = gen.boundary_faces(cells, coords, 1, 0.0)  # edges on y = 0
    corner = np.array([[0]], dtype=np.int64)
    eid = iter(range(1, 10_000))
    rec = lambda conn: np.column_stack([[next(eid) for _ in range(conn.shape[0])], tags[conn]]).astype(np.int64)
    blocks = [(0, 5, 15, rec(corner)), (1, 2, 1, rec(bottom)), (2, 1, 2, rec(cells[:7])), (2, 1, 2, rec(cells[7:]))]
    path = str(tmp_path / "square.msh")
    write_msh41(path, binary, tags[order], coords[order], blocks,
                physical_names=[(0, 30, "corner"), (1, 10, "horizontal"), (2, 20, "domain")],
                entities={0: [(5, [30]), (6, [])], 1: [(2, [10]), (3, [])], 2: [(1, [20])]})
    m = msh.read_msh(path)
    assert m.dim == 2
    rank = np.argsort(np.argsort(tags))  # local id of original node i
    np.testing.assert_array_equal(m.node_tags, np.sort(tags))
    np.testing.assert_array_equal(m.coords[rank], coords)  # bit-exact in both flavours (repr round-trips doubles)
    np.testing.assert_array_equal(m.cells(), rank[cells].astype(np.int32))
    np.testing.assert_array_equal(m.group_elements("horizontal"), rank[bottom].astype(np.int32))
    np.testing.assert_array_equal(m.group_nodes("horizontal"), np.unique(rank[bottom]).astype(np.int32))
    np.testing.assert_array_equal(m.group_nodes("corner"), [rank[0]])
    with pytest.raises(KeyError):
        m.group_nodes("vertical")


def test_malformed_files_are_refused(tmp_path):
    p = tmp_path / "bad.msh"
    p.write_bytes(b"$MeshFormat\n2.2 0 8\n$EndMeshFormat\n")
    with pytest.raises(ValueError, match="version"):
        msh.read_msh(str(p))
    p.write_bytes(b"$MeshFormat\n4.1 0 8\n$EndMeshFormat\n$Elements\n0 0 0 0\n$EndElements\n")
    with pytest.raises(ValueError, match="before"):
        msh.read_msh(str(p))
    p.write_bytes(b"$MeshFormat\n4.1 1 8\n" + struct.pack("<i", 1) + b"\n$EndMeshFormat\n$Nodes\n" + struct.pack("<4Q", 1, 5, 1, 5))
    with pytest.raises(ValueError, match="truncated"):
        msh.read_msh(str(p))


@pytest.mark.skipif(not os.path.exists("/root/reference/meshes/msh/circle_cut.msh"), reason="reference tree not present (GPU box)")
def test_reader_reproduces_the_committed_fixture_of_the_reference_mesh():
    m = msh.read_msh("/root/reference/meshes/msh/circle_cut.msh")
    g = np.load(os.path.join(ROOT, "tests", "golden", "poisson_circle_2D.npz"))
    np.testing.assert_array_equal(m.cells(), g["cells"])
    np.testing.assert_array_equal(m.coords, g["coords"])
    np.testing.assert_array_equal(m.node_tags, g["node_uid"])
    # modules/poisson/inputs: Dirichlet on the group "horizontal"
    np.testing.assert_array_equal(m.group_nodes("horizontal"), g["dirichlet0_nodes"])
