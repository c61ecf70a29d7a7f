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


def test_random_corruption_never_reads_outside_the_relation():
    """Fuzz: thousands of random byte / field mutations of a valid relation.  Every outcome must be an RkError or a successful
    parse whose outputs can be read back — never a crash (the reader bounds-checks every offset it follows)."""
    import vectorsearch_rbac_b200 as vr
    rel, centers, tuples = _make(24, 5, lambda l: (0, 40, 333)[l % 3], late=50)
    rng = np.random.default_rng(123)
    B = pg.BLCKSZ
    nblk = len(rel) // B
    ok = bad = 0
    for it in range(3000):
        b = bytearray(rel)
        for _ in range(int(rng.integers(1, 4))):
            blk = int(rng.integers(0, nblk))
            kind = int(rng.integers(0, 5))
            if kind == 0:                                        # a header field
                off = blk * B + int(rng.choice([12, 14, 16, 18]))
                struct.pack_into("<H", b, off, int(rng.integers(0, 65536)))
            elif kind == 1:                                      # a line pointer
                off = blk * B + pg.HDR + 4 * int(rng.integers(0, 120))
                struct.pack_into("<I", b, off, int(rng.integers(0, 1 << 32)))
            elif kind == 2:                                      # the special space (next block, page id)
                off = blk * B + B - 8 + int(rng.integers(0, 7))
                b[off] = int(rng.integers(0, 256))
            elif kind == 3:                                      # tuple / datum headers somewhere in the item area
                off = blk * B + int(rng.integers(pg.HDR, B - 8))
                b[off] = int(rng.integers(0, 256))
            else:                                                # a burst of random bytes
                off = blk * B + int(rng.integers(0, B - 16))
                b[off:off + 16] = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
        try:
            p = vr.IvfflatPages(bytes(b))
        except vr.RkError:
            bad += 1
            continue
        c, ls, rl, x, tid = p.read()
        assert c.shape == (p.lists, p.dim) and len(tid) == p.tuples == ls[-1] and x.shape == (p.tuples, p.dim)
        p.close()
        ok += 1
    assert bad > 300 and ok > 0


def test_reader_under_address_and_ub_sanitizers(tmp_path):
    """The reader is host code: compile csrc/rk_pgpage.cu alone with g++ -fsanitize=address,undefined (tests/pgpage_asan/
    harness.cpp) and feed it thousands of mutated relations held in exact-size heap blocks — a read one byte outside the relation,
    a misaligned access or an overflow would abort the run."""
    import os
    import shutil
    import subprocess
    gxx = shutil.which("g++")
    if not gxx:
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "harness")
    cmd = [gxx, "-std=c++17", "-g", "-O1", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-fno-omit-frame-pointer",
           "-x", "c++", "-I" + os.path.join(root, "vectorsearch-rbac_b200", "csrc"), "-I" + os.path.join(root, "include"),
           "-I/usr/local/cuda/include", os.path.join(root, "tests", "pgpage_asan", "harness.cpp"), "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0 and "sanitize" in r.stderr and "cannot find" in r.stderr:
        pytest.skip("sanitizer runtime not installed")
    assert r.returncode == 0, r.stderr[-2000:]
    for dim, lists, short in ((24, 5, None), (128, 12, None), (3, 2, False)):
        rel, _, _ = _make(dim, lists, lambda l: (0, 40, 333)[l % 3], short=short, late=30)
        path = tmp_path / f"rel_{dim}.bin"
        path.write_bytes(rel)
        run = subprocess.run([exe, str(path), "4000"], capture_output=True, text=True, env=dict(os.environ, ASAN_OPTIONS="detect_leaks=1"))
        assert run.returncode == 0, (run.stdout[-500:], run.stderr[-3000:])
        ok, bad = [int(x) for x in run.stdout.split()[1::2]]
        assert ok > 0 and bad > 0
