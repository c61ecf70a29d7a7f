"""Test infrastructure: WRITES the main fork of a pgvector ivfflat index the way CREATE INDEX lays it out, so that the reader
(rk_ivfpages_*, csrc/rk_pgpage.cu) has something to read on a box without PostgreSQL.

Restated from the reference (paths relative to /root/reference/pgvector/src):
  ivfutils.c:101-106   IvfflatInitPage: PageInit(page, BLCKSZ, sizeof(IvfflatPageOpaqueData)); nextblkno = Invalid; page_id
  ivfutils.c:135-158   IvfflatAppendPage: new block at the end of the relation, linked through nextblkno
  ivfbuild.c:452-468   CreateMetaPage (block 0: IvfflatMetaPageData after the page header, pd_lower behind it)
  ivfbuild.c:473-517   CreateListPages (from block 1: one MAXALIGNed IvfflatListData per list)
  ivfbuild.c:251-310   InsertTuples (per list: a fresh page, tuples in sort order, itemsz = MAXALIGN(IndexTupleSize))
  ivfbuild.c:240       index_form_tuple: a value of at most 126 data bytes gets the 1-byte varlena header (<= 30 dimensions)
  ivfinsert.c:120-170  a later INSERT appends to the list's insertPage, adding a page at the END of the relation when full
and from PostgreSQL's on-disk format (bufpage.h PageInit / PageAddItem, itemid.h, itup.h, varatt.h; little-endian, BLCKSZ 8192,
MAXALIGN 8), which the reference checkout does not contain.  The reader's parity against a file written by a real server is
therefore UNPINNED; what these tests pin is reader == this restatement of the writer."""
import struct

import numpy as np

BLCKSZ = 8192
INVALID_BLOCK = 0xFFFFFFFF
MAGIC, VERSION, PAGE_ID = 0x14FF1A7, 1, 0xFF84
HDR, SPECIAL = 24, 8


def maxalign(n: int) -> int:
    return (n + 7) & ~7


def new_page() -> bytearray:
    p = bytearray(BLCKSZ)
    struct.pack_into("<HHHH", p, 12, HDR, BLCKSZ - SPECIAL, BLCKSZ - SPECIAL, BLCKSZ | 4)      # pd_lower, pd_upper, pd_special, size|version
    struct.pack_into("<IHH", p, BLCKSZ - SPECIAL, INVALID_BLOCK, 0, PAGE_ID)
    return p


def free_space(p: bytearray) -> int:                       # PageGetFreeSpace
    lower, upper = struct.unpack_from("<HH", p, 12)
    space = upper - lower
    return 0 if space < 4 else space - 4


def add_item(p: bytearray, item: bytes) -> int:            # PageAddItem(..., InvalidOffsetNumber, false, false)
    lower, upper = struct.unpack_from("<HH", p, 12)
    off = (lower - HDR) // 4 + 1
    nlower, nupper = lower + 4, upper - maxalign(len(item))
    assert nlower <= nupper, "page full"
    p[nupper:nupper + len(item)] = item
    struct.pack_into("<I", p, lower, nupper | (1 << 15) | (len(item) << 17))                      # lp_off | LP_NORMAL | lp_len
    struct.pack_into("<HH", p, 12, nlower, nupper)
    return off


def set_next(p: bytearray, blk: int) -> None:
    struct.pack_into("<I", p, BLCKSZ - SPECIAL, blk)


def vector_4b(v: np.ndarray) -> bytes:
    v = np.asarray(v, dtype="<f4")
    return struct.pack("<IhH", (8 + 4 * len(v)) << 2, len(v), 0) + v.tobytes()


def vector_short(v: np.ndarray) -> bytes:
    v = np.asarray(v, dtype="<f4")
    size = 1 + 4 + 4 * len(v)
    assert size <= 0x7F
    return bytes([(size << 1) | 1]) + struct.pack("<hH", len(v), 0) + v.tobytes()


def index_tuple(v: np.ndarray, heap_block: int, heap_off: int, short=None) -> bytes:
    """index_form_tuple for one `vector` key + t_tid.  short=None: as PostgreSQL decides (VARATT_CAN_MAKE_SHORT)."""
    if short is None:
        short = 8 + 4 * len(v) - 4 + 1 <= 0x7F
    datum = vector_short(v) if short else vector_4b(v)
    size = maxalign(8 + len(datum))
    t = struct.pack("<HHHH", (heap_block >> 16) & 0xFFFF, heap_block & 0xFFFF, heap_off, size | 0x4000)
    return (t + datum).ljust(size, b"\0")


def write_index(centers: np.ndarray, lists_tuples, short=None, late_inserts=()) -> bytes:
    """centers [lists, dim]; lists_tuples[l] = [(x, heap_block, heap_off), ...] in stored order;
    late_inserts = [(l, x, heap_block, heap_off), ...] appended after the build as ivfinsert.c does."""
    lists, dim = centers.shape
    pages = [new_page()]
    struct.pack_into("<IIHH", pages[0], HDR, MAGIC, VERSION, dim, lists)
    struct.pack_into("<H", pages[0], 12, HDR + 12)
    # list pages
    pages.append(new_page())
    list_item_size = maxalign(8 + 8 + 4 * dim)
    where = []
    for l in range(lists):
        item = (struct.pack("<II", INVALID_BLOCK, INVALID_BLOCK) + vector_4b(centers[l])).ljust(list_item_size, b"\0")
        if free_space(pages[-1]) < list_item_size:
            pages.append(new_page())
            set_next(pages[-2], len(pages) - 1)
        where.append((len(pages) - 1, add_item(pages[-1], item)))
    insert_page = [0] * lists
    # entry pages
    for l in range(lists):
        pages.append(new_page())
        start = cur = len(pages) - 1
        for x, hb, ho in lists_tuples[l]:
            it = index_tuple(x, hb, ho, short)
            if free_space(pages[cur]) < len(it):
                pages.append(new_page())
                set_next(pages[cur], len(pages) - 1)
                cur = len(pages) - 1
            add_item(pages[cur], it)
        insert_page[l] = cur
        _update_list(pages, where[l], start, cur)
    for l, x, hb, ho in late_inserts:
        it = index_tuple(x, hb, ho, short)
        cur = insert_page[l]
        if free_space(pages[cur]) < len(it):
            pages.append(new_page())
            set_next(pages[cur], len(pages) - 1)
            cur = insert_page[l] = len(pages) - 1
            _update_list(pages, where[l], None, cur)
        add_item(pages[cur], it)
    return b"".join(bytes(p) for p in pages)


def _update_list(pages, where, start, insert) -> None:     # IvfflatUpdateList
    blk, off = where
    lp, = struct.unpack_from("<I", pages[blk], HDR + 4 * (off - 1))
    pos = lp & 0x7FFF
    if start is not None:
        struct.pack_into("<I", pages[blk], pos, start)
    struct.pack_into("<I", pages[blk], pos + 4, insert)
