// rk_pgpage.cu — reader for the on-disk form of a pgvector ivfflat index (the data format on the storage
// side of the path).  Host code only: it runs without a CUDA device and never touches one.
//
// What it restates (paths relative to the reference checkout):
//   pgvector/src/ivfflat.h:33-39   IVFFLAT_VERSION 1, IVFFLAT_MAGIC_NUMBER 0x14FF1A7, IVFFLAT_PAGE_ID 0xFF84,
//                                  IVFFLAT_METAPAGE_BLKNO 0, IVFFLAT_HEAD_BLKNO 1
//   pgvector/src/ivfflat.h:222-246 IvfflatMetaPageData {u32 magic, u32 version, u16 dimensions, u16 lists},
//                                  IvfflatPageOpaqueData {u32 nextblkno, u16 unused, u16 page_id},
//                                  IvfflatListData {u32 startPage, u32 insertPage, Vector center}
//   pgvector/src/vector.h          Vector {int32 vl_len_, int16 dim, int16 unused, float x[]}
//   pgvector/src/ivfbuild.c:452-468 (meta page: contents right after the page header, pd_lower behind them),
//                       :473-517   (list pages: one MAXALIGNed IvfflatListData item per list, chained by nextblkno),
//                       :251-310   (entry pages: index tuples formed by index_form_tuple, one chain per list)
//   pgvector/src/ivfscan.c:36-107  (lists are visited from IVFFLAT_HEAD_BLKNO along nextblkno, items 1..max offset),
//                       :128-168   (a list's tuples: from startPage along nextblkno, items 1..max offset)
// The page, line-pointer, index-tuple and varlena layouts are PostgreSQL's (storage/bufpage.h, storage/itemid.h,
// access/itup.h, varatt.h), little-endian, BLCKSZ 8192, MAXALIGN 8 — PostgreSQL itself is not in the reference checkout, so
// they are restated from its published on-disk format:
//   page header 24 B: pd_lsn 8 | pd_checksum 2 | pd_flags 2 | pd_lower 2 | pd_upper 2 | pd_special 2 | pd_pagesize_version 2 |
//                     pd_prune_xid 4, then 4-byte line pointers {lp_off:15, lp_flags:2, lp_len:15} (LP_NORMAL = 1)
//   index tuple: t_tid {bi_hi u16, bi_lo u16, ip_posid u16} | t_info u16 (size 0x1FFF, has-varwidth 0x4000, has-nulls 0x8000),
//                data at MAXALIGN(8) = 8
//   varlena: 4-byte header (first byte & 3 == 0): total size = le32 >> 2; short header (first byte & 1, byte != 1):
//            total size = byte >> 1, data follows unaligned (index_form_tuple shortens values of up to 126 data bytes:
//            vectors of up to 30 dimensions); compressed (& 3 == 2) and external (byte == 1) values cannot occur for
//            `vector` (STORAGE = external keeps it uncompressed: pgvector/sql/vector.sql:29) and are refused.
#include <cstdint>
#include <cstring>
#include <new>
#include <vector>

#include "rk_host.h"

namespace {

constexpr size_t PG_BLCKSZ = 8192;
constexpr uint32_t PG_INVALID_BLOCK = 0xFFFFFFFFu;
constexpr uint32_t IVF_MAGIC = 0x14FF1A7u, IVF_VERSION = 1u;
constexpr uint16_t IVF_PAGE_ID = 0xFF84u;
constexpr uint32_t PG_PAGE_HEADER = 24, PG_SPECIAL = 8;    // MAXALIGN(sizeof(IvfflatPageOpaqueData))

inline uint16_t rd16(const unsigned char *p) { uint16_t v; memcpy(&v, p, 2); return v; }
inline uint32_t rd32(const unsigned char *p) { uint32_t v; memcpy(&v, p, 4); return v; }

struct Item { const unsigned char *p; uint32_t len; };

}  // namespace

struct rk_ivfpages {
    const unsigned char *rel = nullptr;
    size_t nblocks = 0;
    int32_t dim = 0, lists = 0;
    uint64_t tuples = 0, dead = 0;
    std::vector<const unsigned char *> center;        // [lists] -> float x[dim]
    std::vector<uint32_t> start_page;                 // [lists]
    std::vector<uint32_t> list_start;                 // [lists + 1]
    std::vector<const unsigned char *> tup_x;         // [tuples] -> float x[dim] (possibly unaligned)
    std::vector<uint64_t> tup_tid;                    // [tuples]
};

namespace {

// Page checks of PageIsVerified / PageGetItem territory: a reader of foreign bytes must never index outside the page.
int check_page(const rk_ivfpages &h, uint32_t blk, const unsigned char **out, uint32_t *max_off)
{
    if (blk >= h.nblocks) return rk_fail(RK_ERR_INVALID, "ivfflat pages: block %u beyond the relation (%zu blocks)", blk, h.nblocks);
    const unsigned char *pg = h.rel + (size_t) blk * PG_BLCKSZ;
    const uint32_t lower = rd16(pg + 12), upper = rd16(pg + 14), special = rd16(pg + 16), psv = rd16(pg + 18);
    if (upper == 0) return rk_fail(RK_ERR_INVALID, "ivfflat pages: block %u is a new (zeroed) page", blk);
    if ((psv & 0xFF00u) != PG_BLCKSZ || (psv & 0xFFu) != 4u)
        return rk_fail(RK_ERR_UNSUPPORTED, "ivfflat pages: block %u has page size / layout version 0x%04x (want 8192 / 4)", blk, psv);
    if (lower < PG_PAGE_HEADER || lower > upper || upper > special || special != PG_BLCKSZ - PG_SPECIAL)
        return rk_fail(RK_ERR_INVALID, "ivfflat pages: block %u has a corrupt header (lower %u, upper %u, special %u)", blk, lower, upper, special);
    if (rd16(pg + special + 6) != IVF_PAGE_ID)
        return rk_fail(RK_ERR_INVALID, "ivfflat pages: block %u is not an ivfflat page (page id 0x%04x)", blk, rd16(pg + special + 6));
    *out = pg;
    if (max_off) *max_off = (lower - PG_PAGE_HEADER) / 4;        // PageGetMaxOffsetNumber
    return 0;
}

// 1 = item returned, 0 = line pointer not LP_NORMAL (skipped like a vacuumed slot), < 0 = corrupt
int get_item(const unsigned char *pg, uint32_t blk, uint32_t off /* 1-based */, Item *it)
{
    const uint32_t lp = rd32(pg + PG_PAGE_HEADER + (size_t) (off - 1) * 4);
    const uint32_t lp_off = lp & 0x7FFFu, lp_flags = (lp >> 15) & 3u, lp_len = lp >> 17;
    if (lp_flags != 1u) return 0;
    const uint32_t upper = rd16(pg + 14), special = rd16(pg + 16);
    if (lp_off < upper || lp_off + lp_len > special || (lp_off & 7u))
        return rk_fail(RK_ERR_INVALID, "ivfflat pages: block %u item %u points outside the page (offset %u, length %u)", blk, off, lp_off, lp_len);
    it->p = pg + lp_off;
    it->len = lp_len;
    return 1;
}

// A `vector` datum at d (at most `avail` bytes): returns its float array, checks the dimension.
int vector_datum(const unsigned char *d, uint32_t avail, int32_t want_dim, uint32_t blk, uint32_t off, const unsigned char **x)
{
    if (avail < 1) return rk_fail(RK_ERR_INVALID, "ivfflat pages: block %u item %u holds no datum", blk, off);
    uint32_t size, hdr;
    if (d[0] == 0x01) return rk_fail(RK_ERR_UNSUPPORTED, "ivfflat pages: block %u item %u is an external TOAST pointer", blk, off);
    if (d[0] & 0x01) { size = d[0] >> 1; hdr = 1; }
    else if ((d[0] & 0x03) == 0x02) return rk_fail(RK_ERR_UNSUPPORTED, "ivfflat pages: block %u item %u is a compressed datum", blk, off);
    else {
        if (avail < 4) return rk_fail(RK_ERR_INVALID, "ivfflat pages: block %u item %u is truncated", blk, off);
        size = rd32(d) >> 2; hdr = 4;
    }
    if (size < hdr + 4 || size > avail) return rk_fail(RK_ERR_INVALID, "ivfflat pages: block %u item %u has datum size %u of %u bytes", blk, off, size, avail);
    const int32_t dim = (int16_t) rd16(d + hdr);
    if (dim != want_dim) return rk_fail(RK_ERR_DIM, "different vector dimensions %d and %d", want_dim, dim);       // vector.c:61-67
    if (size != hdr + 4 + 4u * (uint32_t) dim)
        return rk_fail(RK_ERR_INVALID, "ivfflat pages: block %u item %u: %d dimensions do not fit a datum of %u bytes", blk, off, dim, size);
    *x = d + hdr + 4;
    return 0;
}

int parse(rk_ivfpages &h)
{
    const unsigned char *pg;
    int rc = check_page(h, 0, &pg, nullptr);
    if (rc < 0) return rc;
    if (rd16(pg + 12) < PG_PAGE_HEADER + 12) return rk_fail(RK_ERR_INVALID, "ivfflat pages: the meta page holds no meta data");
    const unsigned char *meta = pg + PG_PAGE_HEADER;                 // PageGetContents
    if (rd32(meta) != IVF_MAGIC) return rk_fail(RK_ERR_INVALID, "ivfflat index is not valid");      // ivfutils.c: IvfflatGetMetaPageInfo
    if (rd32(meta + 4) != IVF_VERSION) return rk_fail(RK_ERR_UNSUPPORTED, "ivfflat pages: index version %u (this reader knows %u)", rd32(meta + 4), IVF_VERSION);
    h.dim = rd16(meta + 8);
    h.lists = rd16(meta + 10);
    if (h.dim < 1 || h.dim > 2000) return rk_fail(RK_ERR_INVALID, "ivfflat pages: %d dimensions (1 .. 2000)", h.dim);        // ivfflat.h:24
    if (h.lists < 1) return rk_fail(RK_ERR_INVALID, "ivfflat pages: %d lists", h.lists);
    // list pages (ivfscan.c:45-101)
    std::vector<char> seen(h.nblocks, 0);
    seen[0] = 1;
    for (uint32_t blk = 1; blk != PG_INVALID_BLOCK;) {
        uint32_t maxoff;
        if ((rc = check_page(h, blk, &pg, &maxoff)) < 0) return rc;
        if (seen[blk]) return rk_fail(RK_ERR_INVALID, "ivfflat pages: block %u is linked twice", blk);
        seen[blk] = 1;
        for (uint32_t off = 1; off <= maxoff; off++) {
            Item it;
            if ((rc = get_item(pg, blk, off, &it)) < 0) return rc;
            if (rc == 0) return rk_fail(RK_ERR_INVALID, "ivfflat pages: list page %u has an unused item %u", blk, off);
            if (it.len < 8 + 8) return rk_fail(RK_ERR_INVALID, "ivfflat pages: list item %u of block %u is %u bytes", off, blk, it.len);
            const unsigned char *x;
            if ((rc = vector_datum(it.p + 8, it.len - 8, h.dim, blk, off, &x)) < 0) return rc;
            h.start_page.push_back(rd32(it.p));
            h.center.push_back(x);
        }
        blk = rd32(pg + rd16(pg + 16));
    }
    if ((int64_t) h.center.size() != h.lists)
        return rk_fail(RK_ERR_INVALID, "ivfflat pages: the meta page says %d lists, the list pages hold %zu", h.lists, h.center.size());
    // entry pages (ivfscan.c:128-168), list by list
    h.list_start.assign((size_t) h.lists + 1, 0);
    for (int l = 0; l < h.lists; l++) {
        for (uint32_t blk = h.start_page[(size_t) l]; blk != PG_INVALID_BLOCK;) {
            uint32_t maxoff;
            if ((rc = check_page(h, blk, &pg, &maxoff)) < 0) return rc;
            if (seen[blk]) return rk_fail(RK_ERR_INVALID, "ivfflat pages: block %u is linked twice", blk);
            seen[blk] = 1;
            for (uint32_t off = 1; off <= maxoff; off++) {
                Item it;
                if ((rc = get_item(pg, blk, off, &it)) < 0) return rc;
                if (rc == 0) { h.dead++; continue; }
                if (it.len < 8) return rk_fail(RK_ERR_INVALID, "ivfflat pages: index tuple %u of block %u is %u bytes", off, blk, it.len);
                const uint32_t info = rd16(it.p + 6), tsize = info & 0x1FFFu;
                if (info & 0x8000u) return rk_fail(RK_ERR_INVALID, "ivfflat pages: block %u item %u has a NULL key (ivfflat never stores one)", blk, off);
                if (tsize < 8 || tsize > it.len) return rk_fail(RK_ERR_INVALID, "ivfflat pages: block %u item %u: tuple size %u of %u bytes", blk, off, tsize, it.len);
                const unsigned char *x;
                if ((rc = vector_datum(it.p + 8, tsize - 8, h.dim, blk, off, &x)) < 0) return rc;
                const uint64_t hblk = ((uint64_t) rd16(it.p) << 16) | rd16(it.p + 2);      // BlockIdData {bi_hi, bi_lo}
                h.tup_x.push_back(x);
                h.tup_tid.push_back((hblk << 16) | rd16(it.p + 4));                        // packed like INTEGRATION.md §2
                if (h.tup_x.size() > 0xFFFFFFF0ull) return rk_fail(RK_ERR_UNSUPPORTED, "ivfflat pages: more than 2^32 tuples");
            }
            blk = rd32(pg + rd16(pg + 16));
        }
        h.list_start[(size_t) l + 1] = (uint32_t) h.tup_x.size();
    }
    h.tuples = h.tup_x.size();
    return 0;
}

}  // namespace

extern "C" int rk_ivfpages_open(const void *rel, size_t nbytes, rk_ivfpages **out)
{
    if (!rel || !out) return rk_fail(RK_ERR_INVALID, "rk_ivfpages_open: NULL argument");
    *out = nullptr;
    if (nbytes < 2 * PG_BLCKSZ || nbytes % PG_BLCKSZ)
        return rk_fail(RK_ERR_INVALID, "rk_ivfpages_open: %zu bytes is not a whole number (>= 2) of %zu-byte pages", nbytes, PG_BLCKSZ);
    if (nbytes / PG_BLCKSZ > 0xFFFFFFFEull) return rk_fail(RK_ERR_INVALID, "rk_ivfpages_open: more blocks than a relation can have");
    rk_ivfpages *h = new (std::nothrow) rk_ivfpages;
    if (!h) return rk_fail(RK_ERR_NOMEM, "rk_ivfpages_open: out of memory");
    h->rel = static_cast<const unsigned char *>(rel);
    h->nblocks = nbytes / PG_BLCKSZ;
    int rc;
    try { rc = parse(*h); } catch (const std::bad_alloc &) { rc = rk_fail(RK_ERR_NOMEM, "rk_ivfpages_open: out of memory"); }
    if (rc < 0) { delete h; return rc; }
    *out = h;
    return RK_OK;
}

extern "C" int rk_ivfpages_info(const rk_ivfpages *h, int32_t *dim, int32_t *lists, uint64_t *tuples, uint64_t *unused_items)
{
    if (!h) return rk_fail(RK_ERR_INVALID, "rk_ivfpages_info: NULL handle");
    if (dim) *dim = h->dim;
    if (lists) *lists = h->lists;
    if (tuples) *tuples = h->tuples;
    if (unused_items) *unused_items = h->dead;
    return RK_OK;
}

extern "C" int rk_ivfpages_read(const rk_ivfpages *h, float *centers, uint32_t *list_start, uint32_t *row_list, float *x, uint64_t *tid)
{
    if (!h) return rk_fail(RK_ERR_INVALID, "rk_ivfpages_read: NULL handle");
    const size_t rb = (size_t) h->dim * 4;
    if (centers) for (int l = 0; l < h->lists; l++) memcpy(centers + (size_t) l * h->dim, h->center[(size_t) l], rb);
    if (list_start) memcpy(list_start, h->list_start.data(), ((size_t) h->lists + 1) * 4);
    if (row_list) for (int l = 0; l < h->lists; l++) for (uint32_t i = h->list_start[(size_t) l]; i < h->list_start[(size_t) l + 1]; i++) row_list[i] = (uint32_t) l;
    if (x) for (uint64_t i = 0; i < h->tuples; i++) memcpy(x + i * (size_t) h->dim, h->tup_x[i], rb);
    if (tid) memcpy(tid, h->tup_tid.data(), h->tuples * 8);
    return RK_OK;
}

extern "C" void rk_ivfpages_close(rk_ivfpages *h) { delete h; }
