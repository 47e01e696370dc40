"""The read-only LMDB container reader (fcn_object_detector_b200/lmdb_file.py) against files laid out by the writer below.

The writer is test infrastructure and independent of the reader (it shares no code with it): it lays pages out the way
liblmdb 0.9's mdb.c does -- two meta pages, leaf pages filled from the top down with an ascending slot table, values too
large for a node on overflow pages, branch pages whose first separator key is empty, a stale second meta with a smaller
transaction id.  No liblmdb / py-lmdb exists in this image, so this is a consistency pin of two restatements of the
published layout, not a pin against liblmdb itself; the module header of lmdb_file.py says the same.
"""
import os
import struct

import numpy as np
import pytest

from fcn_object_detector_b200 import lmdb_file
from fcn_object_detector_b200.lmdb_file import LMDBFile, LMDBFormatError


class _Writer(object):
    """Bulk-loads sorted (key, value) pairs into a fresh data.mdb image."""

    def __init__(self, psize=4096):
        self.psize = psize
        self.pages = {}                     # pgno -> bytes (len = psize, or k * psize for an overflow run)
        self.next = 2                       # pages 0 and 1 are the metas
        self.nodemax = (((psize - 16) // 2) & ~1) - 2
        self.counts = dict(branch=0, leaf=0, overflow=0)

    def _alloc(self, n=1):
        p = self.next
        self.next += n
        return p

    def _page(self, pgno, flags, nodes):
        """nodes: list of byte strings (node header + key + data), placed from the end of the page downwards."""
        buf = bytearray(self.psize)
        upper = self.psize
        ptrs = []
        for nd in nodes:
            size = (len(nd) + 1) & ~1
            upper -= size
            buf[upper:upper + len(nd)] = nd
            ptrs.append(upper)
        lower = 16 + 2 * len(ptrs)
        assert lower <= upper, 'page overfull'
        struct.pack_into('<QHHHH', buf, 0, pgno, 0, flags, lower, upper)
        struct.pack_into('<%dH' % len(ptrs), buf, 16, *ptrs)
        self.pages[pgno] = bytes(buf)

    def _leaf_node(self, key, value):
        if 8 + len(key) + len(value) > self.nodemax:
            n = (16 + len(value) + self.psize - 1) // self.psize
            pg = self._alloc(n)
            buf = bytearray(n * self.psize)
            struct.pack_into('<QHHI', buf, 0, pg, 0, 0x04, n)
            buf[16:16 + len(value)] = value
            self.pages[pg] = bytes(buf)
            self.counts['overflow'] += n
            return struct.pack('<HHHH', len(value) & 0xffff, len(value) >> 16, 0x01, len(key)) + key + struct.pack('<Q', pg)
        return struct.pack('<HHHH', len(value) & 0xffff, len(value) >> 16, 0, len(key)) + key + value

    def _fill(self, flags, nodes, first_keys):
        """Packs nodes into as many pages as needed; returns [(first key, pgno)]."""
        out, cur, used, first = [], [], 16, None
        for nd, k in zip(nodes, first_keys):
            size = ((len(nd) + 1) & ~1) + 2
            if cur and used + size > self.psize:
                pg = self._alloc()
                self._page(pg, flags, cur)
                out.append((first, pg))
                cur, used, first = [], 16, None
            if first is None:
                first = k
                if flags == 0x01:               # slot 0 of a branch page: the key is implicit, stored empty
                    nd = nd[:6] + struct.pack('<H', 0) + b''
                    size = 8 + 2
            cur.append(nd)
            used += size
        if cur:
            pg = self._alloc()
            self._page(pg, flags, cur)
            out.append((first, pg))
        self.counts['branch' if flags == 0x01 else 'leaf'] += len(out)
        return out

    def build(self, items):
        items = sorted(items)
        if not items:
            return (1 << 64) - 1, 0
        level = self._fill(0x02, [self._leaf_node(k, v) for k, v in items], [k for k, _ in items])
        depth = 1
        while len(level) > 1:
            nodes = [struct.pack('<HHHH', pg & 0xffff, (pg >> 16) & 0xffff, pg >> 32, len(k)) + k for k, pg in level]
            level = self._fill(0x01, nodes, [k for k, _ in level])
            depth += 1
        return level[0][1], depth

    def meta(self, pgno, txnid, root, depth, entries, db_flags=0, magic=0xBEEFC0DE, version=1):
        buf = bytearray(self.psize)
        struct.pack_into('<QHHHH', buf, 0, pgno, 0, 0x08, 0, 0)
        off = 16
        struct.pack_into('<IIQQ', buf, off, magic, version, 0, 1 << 30)
        off += 24
        struct.pack_into('<IHHQQQQQ', buf, off, self.psize, 0, 0, 0, 0, 0, 0, (1 << 64) - 1)          # free DB: empty
        off += 48
        struct.pack_into('<IHHQQQQQ', buf, off, 0, db_flags, depth, self.counts['branch'], self.counts['leaf'],
                         self.counts['overflow'], entries, root)
        off += 48
        struct.pack_into('<QQ', buf, off, self.next - 1, txnid)
        self.pages[pgno] = bytes(buf)

    def image(self):
        out = bytearray(self.next * self.psize)
        for pg, b in self.pages.items():
            out[pg * self.psize:pg * self.psize + len(b)] = b
        return bytes(out)


def _write_env(directory, items, psize=4096, live_meta=1, **meta_kw):
    """An environment directory as `lmdb.open(directory)` leaves it; the live meta is page `live_meta`, the other one
    is an older transaction that still sees an empty tree."""
    os.makedirs(directory, exist_ok=True)
    w = _Writer(psize)
    root, depth = w.build(items)
    w.meta(1 - live_meta, 1, (1 << 64) - 1, 0, 0)
    w.meta(live_meta, 7, root, depth, len(items), **meta_kw)
    with open(os.path.join(directory, 'data.mdb'), 'wb') as f:
        f.write(w.image())
    return w, depth


def _key(i):
    return '{:0>10d}'.format(i).encode()           # create_training_lmdb.py:161,169,174,184


def test_single_leaf_and_live_meta_choice(tmp_path):
    items = [(_key(i), b'value-%d' % i) for i in range(3)]
    for live in (0, 1):
        d = str(tmp_path / ('env%d' % live))
        _, depth = _write_env(d, items, live_meta=live)
        assert depth == 1
        with LMDBFile(d) as db:
            assert db.page_size == 4096 and db.txnid == 7 and len(db) == 3 and db.depth == 1
            assert [(bytes(k), bytes(v)) for k, v in db] == items
            assert bytes(db.get(_key(1))) == b'value-1' and db.get(b'nope') is None
    # MDB_NOSUBDIR: the path names the file itself
    with LMDBFile(os.path.join(d, 'data.mdb')) as db:
        assert len(list(db)) == 3


def test_empty_environment(tmp_path):
    d = str(tmp_path / 'empty')
    _write_env(d, [])
    with LMDBFile(d) as db:
        assert len(db) == 0 and list(db) == [] and db.get(_key(0)) is None


@pytest.mark.parametrize('psize,n', [(4096, 1500), (512, 1500)])
def test_branch_levels_overflow_values_and_key_order(tmp_path, psize, n):
    """Label-sized records (tens of bytes) and image-sized ones (beyond a node: overflow pages, one spanning several pages)
    mixed in one tree, inserted unsorted; 512-byte pages force a tree at least three levels deep."""
    rs = np.random.RandomState(5)
    items = []
    for i in rs.permutation(n):
        size = int(rs.choice([0, 1, 37, 45, psize // 2 - 30, psize // 2 + 40, psize, 3 * psize + 17]))
        items.append((_key(int(i)), rs.randint(0, 256, size).astype(np.uint8).tobytes()))
    d = str(tmp_path / 'env')
    w, depth = _write_env(d, items, psize=psize)
    assert w.counts['overflow'] > 0 and w.counts['branch'] > 0 and depth >= (3 if psize == 512 else 2)
    with LMDBFile(d) as db:
        got = [(bytes(k), bytes(v)) for k, v in db]
        assert got == sorted(items) and len(db) == n and db.depth == depth
        for k, v in items[::97]:
            assert bytes(db.get(k)) == v
        assert db.get(b'0000000000x') is None and db.get(b'') is None and db.get(b'9' * 10) is None


def test_refusals(tmp_path):
    items = [(_key(i), b'v') for i in range(4)]
    d = str(tmp_path / 'ok')
    _write_env(d, items)
    raw = open(os.path.join(d, 'data.mdb'), 'rb').read()

    def variant(name, data):
        p = tmp_path / name
        p.mkdir()
        (p / 'data.mdb').write_bytes(data)
        return str(p)

    with pytest.raises(FileNotFoundError):
        LMDBFile(str(tmp_path / 'missing'))
    with pytest.raises(LMDBFormatError, match='too short'):
        LMDBFile(variant('short', raw[:600]))
    bad = bytearray(raw)
    struct.pack_into('<I', bad, 16, 0xDEC0EFBE)                     # byte-swapped magic = a big-endian file
    with pytest.raises(LMDBFormatError, match='magic'):
        LMDBFile(variant('magic', bytes(bad)))
    bad = bytearray(raw)
    struct.pack_into('<I', bad, 20, 999)                            # MDB_DEVEL data version
    with pytest.raises(LMDBFormatError, match='version'):
        LMDBFile(variant('version', bytes(bad)))
    with pytest.raises(LMDBFormatError, match='beyond the end'):
        LMDBFile(variant('truncated', raw[:2 * 4096]))
    dd = str(tmp_path / 'dupsort')
    _write_env(dd, items, db_flags=0x04)                            # MDB_DUPSORT
    with pytest.raises(LMDBFormatError, match='flags'):
        LMDBFile(dd)
    # a leaf node that claims to hold a sub-database
    bad = bytearray(raw)
    (ptr0,) = struct.unpack_from('<H', bad, 2 * 4096 + 16)
    struct.pack_into('<H', bad, 2 * 4096 + ptr0 + 4, 0x02)
    with LMDBFile(variant('subdata', bytes(bad))) as db:
        with pytest.raises(LMDBFormatError, match='sub-databases'):
            list(db)
    # a page whose header carries another page's number
    bad = bytearray(raw)
    struct.pack_into('<Q', bad, 2 * 4096, 5)
    with LMDBFile(variant('pgno', bytes(bad))) as db:
        with pytest.raises(LMDBFormatError, match='carries number'):
            list(db)


def test_minibatches_are_the_two_data_layers(tmp_path, built_lib, monkeypatch):
    """models/train_val.prototxt:1-26 + create_training_lmdb.py:172-185: features / labels read in key order, batch_size at
    a time, rewound at the end; image datums stored H x W x 3 un-transposed come back (N,3,H,W)."""
    from fcn_object_detector_b200 import lmdb_datum
    from tests.test_abi_host import _datum_bytes
    monkeypatch.setitem(__import__('sys').modules, 'lmdb', None)    # the package's own reader, whatever the host has
    rs = np.random.RandomState(2)
    n, H, W = 5, 20, 24                                             # 20*24*3 floats = 5.7 KB per record: overflow pages
    imgs = [rs.uniform(0, 1, (H, W, 3)).astype(np.float32) for _ in range(n)]
    labs = [np.asarray([[[3 + i, 4, 10, 12, i % 2]]], np.float64) for i in range(n)]
    f, l = str(tmp_path / 'features'), str(tmp_path / 'labels')
    _write_env(f, [(_key(i), _datum_bytes(imgs[i], packed=bool(i & 1))) for i in range(n)])
    _write_env(l, [(_key(i), _datum_bytes(labs[i])) for i in range(n)])
    assert [k for k, _ in lmdb_datum.read_lmdb(f)] == [_key(i) for i in range(n)]
    assert len(lmdb_datum.read_lmdb(l, limit=2)) == 2
    it = lmdb_datum.minibatches(f, l, batch_size=2)
    order = []
    for _ in range(5):                                              # 10 records = two passes over the 5
        data_in, label = next(it)
        assert data_in.shape == (2, 3, H, W) and label.shape == (2, 1, 1, 5) and data_in.dtype == np.float32
        for j in range(2):
            i = int(label[j, 0, 0, 0]) - 3
            order.append(i)
            assert np.array_equal(data_in[j], imgs[i].transpose(2, 0, 1))
            assert np.array_equal(label[j, 0, 0], labs[i].reshape(5).astype(np.float32))
    assert order == [0, 1, 2, 3, 4, 0, 1, 2, 3, 4]
    assert len(list(lmdb_datum.minibatches(f, l, batch_size=5, epochs=2))) == 2
    # mismatched databases
    l2 = str(tmp_path / 'labels_short')
    _write_env(l2, [(_key(i), _datum_bytes(labs[i])) for i in range(n - 1)])
    with pytest.raises(ValueError, match='different numbers'):
        list(lmdb_datum.minibatches(f, l2, batch_size=1, epochs=1))
    l3 = str(tmp_path / 'labels_keys')
    _write_env(l3, [(_key(i + 1), _datum_bytes(labs[i])) for i in range(n)])
    with pytest.raises(ValueError, match='paired'):
        next(lmdb_datum.minibatches(f, l3, batch_size=1))
    e = str(tmp_path / 'empty')
    _write_env(e, [])
    with pytest.raises(ValueError, match='empty'):
        next(lmdb_datum.minibatches(e, e, batch_size=1))


def test_module_constants_match_the_published_layout():
    assert lmdb_file.MAGIC == 0xBEEFC0DE and lmdb_file.PAGEHDRSZ == 16 and lmdb_file.NODESIZE == 8
    assert lmdb_file._DB.size == 48 and lmdb_file._META.size == 24 and lmdb_file._PAGE.size == 16
