"""Read-only reader for the LMDB container (`data.mdb`) the reference's training data lives in.

/root/reference/scripts/misc/create_training_lmdb.py:108-113,172-185 opens two environments with `lmdb.open(path,
map_size=...)` and puts one record per sample under the key '{:0>10d}'.format(counter); the Data layers of
/root/reference/models/train_val.prototxt:1-26 read them back in key order.  LMDB itself is a third-party dependency of the
reference (py-lmdb / liblmdb 0.9.x, not vendored upstream and not installed in this image), so this module restates the
published on-disk layout of liblmdb 0.9 (`mdb.c`: MDB_meta, MDB_db, MDB_page, MDB_node; 64-bit little-endian build):

  page header, 16 bytes   pgno u64 | pad u16 | flags u16 | lower u16, upper u16  (overflow pages: page count u32)
  page flags              BRANCH 0x01, LEAF 0x02, OVERFLOW 0x04, META 0x08, LEAF2 0x20, SUBP 0x40
  meta, at byte 16 of     magic u32 = 0xBEEFC0DE | version u32 = 1 | address u64 | mapsize u64 | MDB_db free | MDB_db main |
  pages 0 and 1           last_pg u64 | txnid u64;  the meta with the larger txnid is the live one
  MDB_db, 48 bytes        pad u32 (the free DB's holds the page size) | flags u16 | depth u16 | branch_pages u64 |
                          leaf_pages u64 | overflow_pages u64 | entries u64 | root u64 (~0 = empty tree)
  node, 8-byte header     lo u16 | hi u16 | flags u16 | ksize u16 | key | data
                          leaf: data size = lo | hi << 16;  branch: child page = lo | hi << 16 | flags << 32
  node flags              BIGDATA 0x01 (data = u64 number of an overflow page; the value starts 16 bytes into it and runs
                          over `page count` consecutive pages), SUBDATA 0x02, DUPDATA 0x04
  slot table              u16 offsets from byte 16; entries = (lower - 16) / 2; keys ascend (memcmp) within a page and
                          across the in-order walk of the tree

Only what a Caffe dataset uses is read: the unnamed main database, unique keys.  Named sub-databases, duplicate-sorted
databases and big-endian / 32-bit files are refused with an error rather than guessed at.

Not validated against a file written by liblmdb in this image (no liblmdb, no py-lmdb here): tests/test_lmdb_host.py checks
the reader against files laid out by an independent writer in the test that follows the same published layout.  When the
`lmdb` package is importable, `lmdb_datum.read_lmdb` uses it instead of this module.
"""
import mmap
import os
import struct

MAGIC = 0xBEEFC0DE
DATA_VERSION = 1
PAGEHDRSZ = 16
NODESIZE = 8
P_BRANCH, P_LEAF, P_OVERFLOW, P_META, P_LEAF2, P_SUBP = 0x01, 0x02, 0x04, 0x08, 0x20, 0x40
F_BIGDATA, F_SUBDATA, F_DUPDATA = 0x01, 0x02, 0x04
P_INVALID = (1 << 64) - 1
_META = struct.Struct('<IIQQ')                     # magic, version, address, mapsize
_DB = struct.Struct('<IHHQQQQQ')                   # pad, flags, depth, branch, leaf, overflow, entries, root
_PAGE = struct.Struct('<QHHHH')                    # pgno, pad, flags, lower, upper
_NODE = struct.Struct('<HHHH')                     # lo, hi, flags, ksize


class LMDBFormatError(ValueError):
    pass


def _data_file(path):
    """`lmdb.open(path)` makes a directory holding data.mdb; with MDB_NOSUBDIR the path is the file itself."""
    if os.path.isdir(path):
        path = os.path.join(path, 'data.mdb')
    if not os.path.isfile(path):
        raise FileNotFoundError('no LMDB data file at %s' % path)
    return path


class LMDBFile(object):
    """One environment, opened read-only.  Iterating yields (key, value) bytes of the main database in key order."""

    def __init__(self, path):
        self.path = _data_file(path)
        self._f = open(self.path, 'rb')
        size = os.fstat(self._f.fileno()).st_size
        if size < 2 * 512:
            self._f.close()
            raise LMDBFormatError('%s: too short for two meta pages' % self.path)
        self._m = mmap.mmap(self._f.fileno(), 0, access=mmap.ACCESS_READ)
        try:
            first = self._meta(0)
            self.page_size = first['psize']
            if self.page_size < 512 or self.page_size > 65536 or self.page_size & (self.page_size - 1):
                raise LMDBFormatError('%s: page size %d' % (self.path, self.page_size))
            if size < 2 * self.page_size:
                raise LMDBFormatError('%s: shorter than two pages of %d bytes' % (self.path, self.page_size))
            second = self._meta(self.page_size)
            live = second if second['txnid'] > first['txnid'] else first
            self.txnid = live['txnid']
            self.last_page = live['last_pg']
            self.entries = live['main'][6]
            self.depth = live['main'][2]
            self._root = live['main'][7]
            if live['main'][1] & 0x7fff:
                raise LMDBFormatError('%s: main database flags 0x%x (duplicate / integer keys) are not supported' %
                                      (self.path, live['main'][1]))
            if (self.last_page + 1) * self.page_size > size:
                raise LMDBFormatError('%s: last page %d lies beyond the end of the file' % (self.path, self.last_page))
        except Exception:
            self.close()
            raise

    def _meta(self, off):
        pgno, _, flags, _, _ = _PAGE.unpack_from(self._m, off)
        if not flags & P_META:
            raise LMDBFormatError('%s: page at byte %d is not a meta page' % (self.path, off))
        magic, version, _, mapsize = _META.unpack_from(self._m, off + PAGEHDRSZ)
        if magic != MAGIC:
            raise LMDBFormatError('%s: magic 0x%08x is not an LMDB file (or not a little-endian one)' % (self.path, magic))
        if version != DATA_VERSION:
            raise LMDBFormatError('%s: data format version %d, reader knows %d' % (self.path, version, DATA_VERSION))
        free = _DB.unpack_from(self._m, off + PAGEHDRSZ + _META.size)
        main = _DB.unpack_from(self._m, off + PAGEHDRSZ + _META.size + _DB.size)
        last_pg, txnid = struct.unpack_from('<QQ', self._m, off + PAGEHDRSZ + _META.size + 2 * _DB.size)
        return {'psize': free[0], 'free': free, 'main': main, 'last_pg': last_pg, 'txnid': txnid, 'mapsize': mapsize}

    # -- pages -----------------------------------------------------------------------------------------------------------
    def _page(self, pgno):
        if pgno > self.last_page:
            raise LMDBFormatError('%s: page %d beyond last page %d' % (self.path, pgno, self.last_page))
        off = pgno * self.page_size
        got, _, flags, lower, upper = _PAGE.unpack_from(self._m, off)
        if got != pgno:
            raise LMDBFormatError('%s: page %d carries number %d' % (self.path, pgno, got))
        return off, flags, lower, upper

    def _nodes(self, off, lower):
        n = (lower - PAGEHDRSZ) >> 1
        if lower < PAGEHDRSZ or lower > self.page_size:
            raise LMDBFormatError('%s: slot table bound %d' % (self.path, lower))
        return struct.unpack_from('<%dH' % n, self._m, off + PAGEHDRSZ)

    def _value(self, node_off, lo, hi, flags, ksize):
        size = lo | (hi << 16)
        at = node_off + NODESIZE + ksize
        if flags & (F_SUBDATA | F_DUPDATA):
            raise LMDBFormatError('%s: named sub-databases / duplicate keys are not supported' % self.path)
        if not flags & F_BIGDATA:
            return self._m[at:at + size]
        (pgno,) = struct.unpack_from('<Q', self._m, at)
        off, pflags, _, _ = self._page(pgno)
        if not pflags & P_OVERFLOW:
            raise LMDBFormatError('%s: page %d is not an overflow page' % (self.path, pgno))
        (count,) = struct.unpack_from('<I', self._m, off + 12)
        if PAGEHDRSZ + size > count * self.page_size or pgno + count - 1 > self.last_page:
            raise LMDBFormatError('%s: value of %d bytes does not fit the %d overflow pages at %d' % (self.path, size, count, pgno))
        return self._m[off + PAGEHDRSZ:off + PAGEHDRSZ + size]

    def _walk(self, pgno, depth):
        if depth > 64:
            raise LMDBFormatError('%s: tree deeper than 64 levels (a cycle?)' % self.path)
        off, flags, lower, _ = self._page(pgno)
        if flags & (P_LEAF2 | P_SUBP):
            raise LMDBFormatError('%s: fixed-size duplicate pages are not supported' % self.path)
        if flags & P_LEAF:
            for p in self._nodes(off, lower):
                lo, hi, nflags, ksize = _NODE.unpack_from(self._m, off + p)
                key = self._m[off + p + NODESIZE:off + p + NODESIZE + ksize]
                yield key, self._value(off + p, lo, hi, nflags, ksize)
        elif flags & P_BRANCH:
            for p in self._nodes(off, lower):
                lo, hi, nflags, _ = _NODE.unpack_from(self._m, off + p)
                for kv in self._walk(lo | (hi << 16) | (nflags << 32), depth + 1):
                    yield kv
        else:
            raise LMDBFormatError('%s: page %d has flags 0x%x, neither branch nor leaf' % (self.path, pgno, flags))

    # -- public -----------------------------------------------------------------------------------------------------------
    def __iter__(self):
        if self._root == P_INVALID:
            return iter(())
        return self._walk(self._root, 0)

    def __len__(self):
        return int(self.entries)

    def get(self, key, default=None):
        """Value stored under `key` (bytes), by descent from the root: the last child whose separator key <= key."""
        key = bytes(key)
        if self._root == P_INVALID:
            return default
        pgno = self._root
        for _ in range(65):
            off, flags, lower, _ = self._page(pgno)
            ptrs = self._nodes(off, lower)
            if flags & P_BRANCH:
                child = None
                for i, p in enumerate(ptrs):
                    lo, hi, nflags, ksize = _NODE.unpack_from(self._m, off + p)
                    sep = self._m[off + p + NODESIZE:off + p + NODESIZE + ksize]
                    if i and sep > key:                 # slot 0 of a branch page has an implicit (empty) key
                        break
                    child = lo | (hi << 16) | (nflags << 32)
                if child is None:
                    raise LMDBFormatError('%s: empty branch page %d' % (self.path, pgno))
                pgno = child
                continue
            if not flags & P_LEAF:
                raise LMDBFormatError('%s: page %d has flags 0x%x' % (self.path, pgno, flags))
            for p in ptrs:
                lo, hi, nflags, ksize = _NODE.unpack_from(self._m, off + p)
                if self._m[off + p + NODESIZE:off + p + NODESIZE + ksize] == key:
                    return self._value(off + p, lo, hi, nflags, ksize)
            return default
        raise LMDBFormatError('%s: tree deeper than 64 levels (a cycle?)' % self.path)

    def close(self):
        m, f = getattr(self, '_m', None), getattr(self, '_f', None)
        self._m = self._f = None
        if m is not None:
            m.close()
        if f is not None:
            f.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
