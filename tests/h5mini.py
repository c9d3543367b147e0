"""Minimal libhdf5-free HDF5 reader (numpy only) for the reference's field dumps.  TEST INFRASTRUCTURE.

Reads exactly the subset `saveRankData` produces (HDF5_output.cpp:10-79; physical layout in SURVEY.md section 8b):
superblock v0, old-style groups (v1 B-tree `TREE` + local heap `HEAP` + symbol nodes `SNOD`), v1 object headers
with continuation blocks, v1 attribute messages holding int32 scalars, and contiguous (layout v3 class 1) datasets of
the 48-byte compound {Ex,Ey,Ez,Bx,By,Bz}.  Used to (a) turn the shipped dumps under
/root/reference/Simulation/Fields into the committed golden fixtures (tests/golden/make_golden.py) and (b) check the
files written by the repo's own libhdf5-free writer.
"""
import mmap
import struct
import numpy as np

GRID_DTYPE = np.dtype([(n, "<f8") for n in ("Ex", "Ey", "Ez", "Bx", "By", "Bz")])


class H5Mini:
    def __init__(self, path):
        with open(path, "rb") as f:                          # mapped, not read: the 65 536-tile dump of an 8192^2 grid is 4 GB
            self.b = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ)
        b = self.b
        if b[:8] != b"\x89HDF\r\n\x1a\n":
            raise ValueError("not an HDF5 file")
        if b[8] != 0 or b[13] != 8 or b[14] != 8:
            raise ValueError("only superblock v0 with 8-byte offsets/lengths is supported")
        self.leaf_k, self.internal_k = struct.unpack_from("<HH", b, 16)
        self.eof = struct.unpack_from("<Q", b, 40)[0]
        _, self.root_ohdr, cache, _ = struct.unpack_from("<QQII", b, 56)
        if cache == 1:
            self.root_btree, self.root_heap = struct.unpack_from("<QQ", b, 80)
        else:
            self.root_btree, self.root_heap = self._stab(self.root_ohdr)

    # -- object headers -------------------------------------------------------------------------------------
    def messages(self, addr):
        """Yield (type, flags, payload bytes) of a v1 object header, following continuation messages."""
        b = self.b
        ver, _, nmsgs, _ref, hsize = struct.unpack_from("<BBHII", b, addr)
        if ver != 1:
            raise ValueError("only v1 object headers are supported")
        blocks = [(addr + 16, hsize)]
        seen = 0
        while blocks and seen < nmsgs:
            pos, size = blocks.pop(0)
            end = pos + size
            while pos + 8 <= end and seen < nmsgs:
                mtype, msize, flags = struct.unpack_from("<HHB", b, pos)
                data = b[pos + 8: pos + 8 + msize]
                pos += 8 + msize
                seen += 1
                if mtype == 0x0010:
                    off, length = struct.unpack_from("<QQ", data, 0)
                    blocks.append((off, length))
                else:
                    yield mtype, flags, data

    def _stab(self, ohdr):
        for mtype, _, data in self.messages(ohdr):
            if mtype == 0x0011:
                return struct.unpack_from("<QQ", data, 0)
        raise ValueError("object at %d is not an old-style group" % ohdr)

    # -- groups ---------------------------------------------------------------------------------------------
    def _heap_data(self, heap):
        if self.b[heap:heap + 4] != b"HEAP":
            raise ValueError("bad local heap")
        return struct.unpack_from("<Q", self.b, heap + 24)[0]

    def _walk(self, node, heap_data, out):
        b = self.b
        if b[node:node + 4] == b"TREE":
            _ntype, level, used = struct.unpack_from("<BBH", b, node + 4)
            for k in range(used):
                child = struct.unpack_from("<Q", b, node + 24 + 8 + 16 * k)[0]
                self._walk(child, heap_data, out)
        elif b[node:node + 4] == b"SNOD":
            nsym = struct.unpack_from("<H", b, node + 6)[0]
            for k in range(nsym):
                noff, ohdr = struct.unpack_from("<QQ", b, node + 8 + 40 * k)
                s = heap_data + noff
                name = b[s:b.find(b"\0", s)].decode()
                out.append((name, ohdr))
        else:
            raise ValueError("unexpected node signature at %d" % node)

    def links(self, ohdr=None):
        """[(name, object header address)] of a group in B-tree (name-sorted) order."""
        btree, heap = (self.root_btree, self.root_heap) if ohdr is None else self._stab(ohdr)
        out = []
        self._walk(btree, self._heap_data(heap), out)
        return out

    # -- attributes / datasets --------------------------------------------------------------------------------
    def attrs(self, ohdr):
        out = {}
        for mtype, _, data in self.messages(ohdr):
            if mtype != 0x000C:
                continue
            ver, _, nsize, tsize, ssize = struct.unpack_from("<BBHHH", data, 0)
            if ver != 1:
                raise ValueError("only v1 attribute messages are supported")
            pad = lambda n: (n + 7) & ~7
            name = data[8:8 + nsize].split(b"\0")[0].decode()
            voff = 8 + pad(nsize) + pad(tsize) + pad(ssize)
            out[name] = struct.unpack_from("<i", data, voff)[0]
        return out

    def dataset(self, ohdr):
        dims = addr = size = None
        for mtype, _, data in self.messages(ohdr):
            if mtype == 0x0001:
                rank = data[1]
                dims = struct.unpack_from("<%dQ" % rank, data, 8)
            elif mtype == 0x0008:
                if data[0] != 3 or data[1] != 1:
                    raise ValueError("only contiguous layout v3 is supported")
                addr, size = struct.unpack_from("<QQ", data, 2)
        n = int(np.prod(dims))
        if size != n * GRID_DTYPE.itemsize:
            raise ValueError("unexpected element size")
        return np.frombuffer(self.b, dtype=GRID_DTYPE, count=n, offset=addr).reshape(dims)

    def tiles(self):
        """{gid: (attrs dict, (TY,TX) structured array incl. guards)} for every /Tile_<gid> group."""
        out = {}
        for name, ohdr in self.links():
            if not name.startswith("Tile_"):
                continue
            sub = dict(self.links(ohdr))
            out[int(name[5:])] = (self.attrs(ohdr), self.dataset(sub["fields"]))
        return out


def assemble(paths, nx, ny, tnx, tny, guard):
    """Rebuild the six global (ny,nx) planes from per-rank dump files, placing interiors exactly as the reference's
    consumer does (File_reader.py:111-117).  Returns (planes dict, owner dict gid->currentRank, tiles dict)."""
    planes = {n: np.zeros((ny, nx)) for n in GRID_DTYPE.names}
    owner, tiles = {}, {}
    for p in paths:
        for gid, (a, d) in H5Mini(p).tiles().items():
            r0, c0 = a["tileRow"] * tny, a["tileCol"] * tnx
            for n in GRID_DTYPE.names:
                planes[n][r0:r0 + tny, c0:c0 + tnx] = d[n][guard:guard + tny, guard:guard + tnx]
            owner[gid] = a["currentRank"]
            tiles[gid] = (a, d)
    return planes, owner, tiles
