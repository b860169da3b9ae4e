"""Gmsh MSH 4.1 reader (binary and ASCII), host side, numpy only.

Stands in for Arcane's mesh reader on the reference's own case meshes (meshes/msh/*.msh are MSH 4.1 binary): it
delivers what the assembly path reads from Arcane — cellNode() connectivity of the P1 cells, node coordinates, node
unique ids (Arcane keeps the Gmsh node tag as uniqueId, which is what the reference's result files are keyed on,
femutils/FemUtils.cc:100-170) — plus the named groups the case files refer to (<surface>horizontal</surface>,
<node>topLeftCorner</node>): node groups and boundary-face groups built from the physical names.

File layout followed: $MeshFormat / $PhysicalNames / $Entities / $Nodes / $Elements of the published MSH 4.1
specification (size_t = 8 bytes, int = 4 bytes, double = 8 bytes, native little endian when the header's "one" reads 1).
"""
import io
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import numpy as np

# element type -> (dimension, nodes); P1 simplices + points are what the assembly path uses
ELEMENT_TYPES = {15: (0, 1), 1: (1, 2), 2: (2, 3), 4: (3, 4), 3: (2, 4), 5: (3, 8), 6: (3, 6), 7: (3, 5), 8: (1, 3), 9: (2, 6), 11: (3, 10)}


@dataclass
class MshMesh:
    node_tags: np.ndarray  # int64 [nb_node]  (Arcane uniqueId)
    coords: np.ndarray  # float64 [nb_node, 3]
    # per element type: (entity_dim, entity_tag int32 [n], element_tag int64 [n], node index int32 [n, nodes]) — node
    # indices are positions in node_tags/coords (local ids), not Gmsh tags
    elements: Dict[int, Tuple[int, np.ndarray, np.ndarray, np.ndarray]] = field(default_factory=dict)
    physical_names: Dict[str, Tuple[int, int]] = field(default_factory=dict)  # name -> (dim, physical tag)
    entity_physicals: Dict[Tuple[int, int], List[int]] = field(default_factory=dict)  # (dim, entity tag) -> physical tags

    @property
    def dim(self):
        return 3 if 4 in self.elements else 2

    def cells(self):
        """P1 cells of the highest dimension present: Tetraedron4 (type 4) or Triangle3 (type 2), int32 [nb_cell, dim+1]."""
        t = 4 if 4 in self.elements else 2
        if t not in self.elements:
            raise ValueError("the mesh holds neither Tetraedron4 nor Triangle3 cells")
        return np.ascontiguousarray(self.elements[t][3], dtype=np.int32)

    def _entities_of(self, name):
        if name not in self.physical_names:
            raise KeyError(f"no physical group named {name!r} (have {sorted(self.physical_names)})")
        dim, tag = self.physical_names[name]
        ents = {e for (d, e), phys in self.entity_physicals.items() if d == dim and tag in phys}
        return dim, ents

    def group_elements(self, name):
        """Node indices [n, nodes] of the elements belonging to the physical group (faces of a <surface>, points of a <node>)."""
        dim, ents = self._entities_of(name)
        out = []
        for etype, (edim, etag, _, conn) in self.elements.items():
            if edim != dim:
                continue
            mask = np.isin(etag, list(ents))
            if mask.any():
                out.append(conn[mask])
        if not out:
            return np.zeros((0, max(dim + 1, 1)), dtype=np.int32)
        if len({o.shape[1] for o in out}) != 1:
            raise ValueError(f"group {name!r} mixes element types")
        return np.ascontiguousarray(np.concatenate(out, axis=0), dtype=np.int32)

    def group_nodes(self, name):
        """Sorted local node ids of the group (what the reference's `group.nodeGroup()` enumerates)."""
        return np.unique(self.group_elements(name).ravel()).astype(np.int32)


class _Reader:
    def __init__(self, data):
        self.b = io.BytesIO(data)

    def line(self):
        return self.b.readline().decode("latin-1").strip()

    def expect(self, text):
        got = self.line()
        while got == "":
            got = self.line()
        if got != text:
            raise ValueError(f"MSH: expected {text!r}, found {got!r}")

    def raw(self, fmt, count=1):
        dt = np.dtype(fmt)
        size = dt.itemsize * count
        buf = self.b.read(size)
        if len(buf) != size:
            raise ValueError("MSH: truncated binary section")
        return np.frombuffer(buf, dtype=dt, count=count)

    def skip_section(self, name):
        end = f"$End{name}".encode()
        rest = self.b.read()
        i = rest.find(end)
        if i < 0:
            raise ValueError(f"MSH: missing $End{name}")
        self.b = io.BytesIO(rest[i + len(end):])
        self.b.readline()


def read_msh(path):
    data = open(path, "rb").read()
    r = _Reader(data)
    r.expect("$MeshFormat")
    version, file_type, data_size = r.line().split()
    if not version.startswith("4.1"):
        raise ValueError(f"MSH: version {version} not supported (4.1 expected)")
    binary = int(file_type) == 1
    if int(data_size) != 8:
        raise ValueError("MSH: data size must be 8")
    if binary:
        one = r.raw("<i4")[0]
        if one != 1:
            raise ValueError("MSH: big-endian binary files are not supported")
        r.b.readline()
    r.expect("$EndMeshFormat")

    mesh = MshMesh(np.zeros(0, np.int64), np.zeros((0, 3)))
    tag_to_index = None
    while True:
        header = r.line()
        if header == "":
            if r.b.tell() >= len(r.b.getbuffer()):
                break
            continue
        name = header[1:]
        if name == "PhysicalNames":
            n = int(r.line())
            for _ in range(n):
                parts = r.line().split(" ", 2)
                mesh.physical_names[parts[2].strip().strip('"')] = (int(parts[0]), int(parts[1]))
            r.expect("$EndPhysicalNames")
        elif name == "Entities":
            _read_entities(r, mesh, binary)
            r.expect("$EndEntities")
        elif name == "Nodes":
            tag_to_index = _read_nodes(r, mesh, binary)
            r.expect("$EndNodes")
        elif name == "Elements":
            if tag_to_index is None:
                raise ValueError("MSH: $Elements before $Nodes")
            _read_elements(r, mesh, binary, tag_to_index)
            r.expect("$EndElements")
        else:
            r.skip_section(name)
    return mesh


def _read_entities(r, mesh, binary):
    if binary:
        counts = r.raw("<u8", 4)
    else:
        counts = [int(x) for x in r.line().split()]
    for dim in range(4):
        for _ in range(int(counts[dim])):
            if binary:
                tag = int(r.raw("<i4")[0])
                r.raw("<f8", 3 if dim == 0 else 6)
                nphys = int(r.raw("<u8")[0])
                phys = [int(x) for x in r.raw("<i4", nphys)]
                if dim > 0:
                    nb = int(r.raw("<u8")[0])
                    r.raw("<i4", nb)
            else:
                f = r.line().split()
                tag = int(f[0])
                o = 4 if dim == 0 else 7
                nphys = int(f[o])
                phys = [int(x) for x in f[o + 1:o + 1 + nphys]]
            mesh.entity_physicals[(dim, tag)] = phys


def _read_nodes(r, mesh, binary):
    if binary:
        nblocks, nnodes, _, _ = (int(x) for x in r.raw("<u8", 4))
    else:
        nblocks, nnodes, _, _ = (int(x) for x in r.line().split())
    tags, xyz = [], []
    for _ in range(nblocks):
        if binary:
            _, _, parametric = (int(x) for x in r.raw("<i4", 3))
            n = int(r.raw("<u8")[0])
            t = r.raw("<u8", n).astype(np.int64)
            if parametric:
                raise ValueError("MSH: parametric node blocks are not supported")
            c = r.raw("<f8", 3 * n).reshape(n, 3)
        else:
            f = r.line().split()
            parametric, n = int(f[2]), int(f[3])
            if parametric:
                raise ValueError("MSH: parametric node blocks are not supported")
            t = np.array([int(r.line()) for _ in range(n)], dtype=np.int64)
            c = np.array([[float(x) for x in r.line().split()[:3]] for _ in range(n)], dtype=np.float64).reshape(n, 3)
        tags.append(t)
        xyz.append(c)
    if binary:
        r.b.readline()
    tags = np.concatenate(tags) if tags else np.zeros(0, np.int64)
    xyz = np.concatenate(xyz) if xyz else np.zeros((0, 3))
    if tags.shape[0] != nnodes:
        raise ValueError("MSH: node count mismatch")
    order = np.argsort(tags, kind="stable")  # local ids in increasing unique id, like a sequential Arcane read
    mesh.node_tags = tags[order]
    mesh.coords = np.ascontiguousarray(xyz[order])
    return {int(t): i for i, t in enumerate(mesh.node_tags)}


def _read_elements(r, mesh, binary, tag_to_index):
    if binary:
        nblocks, nelem, _, _ = (int(x) for x in r.raw("<u8", 4))
    else:
        nblocks, nelem, _, _ = (int(x) for x in r.line().split())
    lut_max = int(mesh.node_tags.max()) if mesh.node_tags.size else 0
    lut = np.full(lut_max + 1, -1, dtype=np.int64)
    lut[mesh.node_tags] = np.arange(mesh.node_tags.shape[0])
    acc: Dict[int, List] = {}
    total = 0
    for _ in range(nblocks):
        if binary:
            edim, etag, etype = (int(x) for x in r.raw("<i4", 3))
            n = int(r.raw("<u8")[0])
        else:
            f = r.line().split()
            edim, etag, etype, n = int(f[0]), int(f[1]), int(f[2]), int(f[3])
        if etype not in ELEMENT_TYPES:
            raise ValueError(f"MSH: element type {etype} not supported")
        nn = ELEMENT_TYPES[etype][1]
        if binary:
            rec = r.raw("<u8", n * (1 + nn)).reshape(n, 1 + nn).astype(np.int64)
        else:
            rec = np.array([[int(x) for x in r.line().split()] for _ in range(n)], dtype=np.int64).reshape(n, 1 + nn)
        conn = lut[rec[:, 1:]]
        if (conn < 0).any():
            raise ValueError("MSH: element refers to an unknown node tag")
        acc.setdefault(etype, []).append((edim, np.full(n, etag, dtype=np.int32), rec[:, 0], conn.astype(np.int32)))
        total += n
    if binary:
        r.b.readline()
    if total != nelem:
        raise ValueError("MSH: element count mismatch")
    for etype, blocks in acc.items():
        edim = blocks[0][0]
        mesh.elements[etype] = (edim, np.concatenate([b[1] for b in blocks]), np.concatenate([b[2] for b in blocks]),
                                np.concatenate([b[3] for b in blocks], axis=0))
