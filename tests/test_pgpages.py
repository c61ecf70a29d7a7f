"""The ivfflat index-file reader (rk_ivfpages_*) against tests/pgpages.py, which writes the relation the way
pgvector's CREATE INDEX / INSERT lay it out (ivfbuild.c:251-310, 452-517; ivfinsert.c:120-170).  Host code: no GPU needed.
Parity against a file written by a real PostgreSQL server is unpinned (none can run here); these tests pin the reader to the
restated writer and to the error behaviour on corrupt pages."""
import struct

import numpy as np
import pytest

from tests import pgpages as pg


def _make(dim, lists, per_list, seed=0, short=None, late=0):
    rng = np.random.default_rng(seed)
    centers = rng.standard_normal((lists, dim)).astype(np.float32)
    tuples, n = [], 0
    for l in range(lists):
        cnt = per_list(l)
        x = (centers[l] + 0.1 * rng.standard_normal((cnt, dim))).astype(np.float32)
        tuples.append([(x[i], (n + i) // 200 + (70000 if l % 2 else 0), (n + i) % 200 + 1) for i in range(cnt)])
        n += cnt
    late_inserts = []
    for j in range(late):
        l = int(rng.integers(0, lists))
        late_inserts.append((l, rng.standard_normal(dim).astype(np.float32), 900000 + j, 1 + j % 50))
    rel = pg.write_index(centers, tuples, short=short, late_inserts=late_inserts)
    for l, x, hb, ho in late_inserts:
        tuples[l].append((x, hb, ho))
    return rel, centers, tuples


def _check(rel, centers, tuples):
    import vectorsearch_rbac_b200 as vr
    p = vr.IvfflatPages(rel)
    lists, dim = centers.shape
    assert (p.dim, p.lists, p.tuples, p.unused_items) == (dim, lists, sum(len(t) for t in tuples), 0)
    c, ls, rl, x, tid = p.read()
    np.testing.assert_array_equal(c.view(np.uint32), centers.view(np.uint32))
    np.testing.assert_array_equal(ls, np.cumsum([0] + [len(t) for t in tuples]))
    flat = [t for l in tuples for t in l]
    if flat:
        np.testing.assert_array_equal(x.view(np.uint32), np.stack([t[0] for t in flat]).view(np.uint32))
    np.testing.assert_array_equal(tid, np.array([(t[1] << 16) | t[2] for t in flat], dtype=np.uint64))
    np.testing.assert_array_equal(rl, np.repeat(np.arange(lists), [len(t) for t in tuples]))
    p.close()


@pytest.mark.parametrize("dim,lists,short", [(3, 1, None), (3, 5, False), (30, 7, None), (31, 7, None), (128, 300, None), (2000, 3, None)])
def test_reader_returns_what_create_index_wrote(dim, lists, short):
    # empty lists, lists of one page and of several pages; 128 x 300: 15 list items per page -> 20 list pages;
    # <= 30 dimensions: 1-byte varlena headers (unaligned floats), else 4-byte headers
    rel, centers, tuples = _make(dim, lists, lambda l: (0, 1, 37, 700)[l % 4] if dim < 2000 else 5, short=short)
    assert len(rel) % pg.BLCKSZ == 0
    _check(rel, centers, tuples)


def test_reader_follows_pages_added_by_later_inserts():
    # ivfinsert.c: tuples inserted after the build go to the list's insertPage; a full page gets a successor at the END of the
    # relation, so a list's chain is not contiguous
    rel, centers, tuples = _make(64, 6, lambda l: 150, late=400)
    _check(rel, centers, tuples)


def test_unused_line_pointers_are_skipped_and_counted():
    import vectorsearch_rbac_b200 as vr
    rel, centers, tuples = _make(16, 2, lambda l: 10)
    b = bytearray(rel)
    first_entry = 2 * pg.BLCKSZ                                   # block 0 meta, 1 list page, 2 = list 0's first page
    lp, = struct.unpack_from("<I", b, first_entry + pg.HDR + 4 * 3)
    struct.pack_into("<I", b, first_entry + pg.HDR + 4 * 3, lp & ~(3 << 15))        # item 4 -> LP_UNUSED
    p = vr.IvfflatPages(bytes(b))
    assert (p.tuples, p.unused_items) == (19, 1)
    _, ls, _, x, tid = p.read()
    np.testing.assert_array_equal(ls, [0, 9, 19])
    exp = [t for i, t in enumerate(tuples[0]) if i != 3] + tuples[1]
    np.testing.assert_array_equal(tid, np.array([(t[1] << 16) | t[2] for t in exp], dtype=np.uint64))
    np.testing.assert_array_equal(x, np.stack([t[0] for t in exp]))


def _corrupt(rel, off, fmt, *vals):
    b = bytearray(rel)
    struct.pack_into(fmt, b, off, *vals)
    return bytes(b)


def test_corrupt_or_foreign_pages_are_errors_not_wild_reads():
    import vectorsearch_rbac_b200 as vr
    rel, centers, tuples = _make(16, 3, lambda l: 300)             # 300 tuples of 80 bytes: several pages per list
    B = pg.BLCKSZ
    lp0, = struct.unpack_from("<I", rel, 2 * B + pg.HDR)
    item0 = 2 * B + (lp0 & 0x7FFF)
    cases = [
        (rel[:-100], "whole number"),
        (rel[:B], "whole number"),
        (_corrupt(rel, pg.HDR, "<I", 0x12345678), "ivfflat index is not valid"),                    # ivfutils.c magic check
        (_corrupt(rel, pg.HDR + 4, "<I", 2), "index version 2"),
        (_corrupt(rel, B + B - 2, "<H", 0xFF80), "not an ivfflat page"),                            # a btree-ish page id
        (_corrupt(rel, 2 * B + 18, "<H", 4096 | 4), "page size"),
        (_corrupt(rel, 2 * B + 14, "<H", 8190), "corrupt header"),                                  # pd_upper > pd_special
        (_corrupt(rel, 2 * B + pg.HDR, "<I", 8000 | (1 << 15) | (400 << 17)), "outside the page"),  # item runs into the special space
        (_corrupt(rel, 2 * B + B - 8, "<I", 2), "linked twice"),                                    # nextblkno -> itself
        (_corrupt(rel, 2 * B + B - 8, "<I", 100000), "beyond the relation"),
        (_corrupt(rel, item0 + 6, "<H", 80 | 0xC000), "NULL key"),
        (_corrupt(rel, item0 + 6, "<H", 800 | 0x4000), "tuple size"),
        (_corrupt(rel, item0 + 9, "<h", 15), "different vector dimensions 16 and 15"),              # vector.c:61-67 wording
        (_corrupt(rel, item0 + 8, "<B", 0x01), "external TOAST"),
        (_corrupt(rel, pg.HDR + 10, "<H", 4), "the meta page says 4 lists"),
        (rel[:2 * B] + bytes(B) + rel[3 * B:], "new (zeroed) page"),
    ]
    for bad, msg in cases:
        with pytest.raises(vr.RkError) as e:
            vr.IvfflatPages(bad)
        assert msg in str(e.value), (msg, str(e.value))
    vr.IvfflatPages(rel).close()                                    # the untouched file still reads
