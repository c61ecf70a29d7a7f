/* A plain C99 client of include/rbacknn.h — what a pgvector-side shim (or the reference's C++ benches through
 * extern "C") would compile.  Mode "nodevice": the no-GPU error path.  Mode "scan": a tiny filtered scan whose answers
 * the caller (tests/test_c_client.py) checks against the oracle. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rbacknn.h"

static int fail(const char *what)
{
    fprintf(stderr, "%s: %s\n", what, rk_last_error());
    return 1;
}

int main(int argc, char **argv)
{
    const char *mode = argc > 1 ? argv[1] : "nodevice";
    rk_config cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.abi_version = RK_ABI_VERSION;
    cfg.dim = 4;
    cfg.metric = RK_METRIC_L2;
    cfg.device = 0;
    cfg.k_max = 8;
    if (rk_abi_version() != RK_ABI_VERSION) { fprintf(stderr, "ABI mismatch\n"); return 1; }
    rk_index *ix = rk_open(&cfg);
    if (!strcmp(mode, "nodevice")) {
        if (ix) { fprintf(stderr, "rk_open succeeded without a device?\n"); rk_close(ix); return 1; }
        printf("rk_open failed as expected: %s\n", rk_last_error());
        /* argument validation needs no device either */
        if (rk_finalize(NULL) != RK_ERR_INVALID) return 1;
        if (rk_gettuple(NULL, NULL, NULL, NULL) != RK_ERR_INVALID) return 1;
        return 0;
    }
    if (!ix) return fail("rk_open");
    /* six rows in three documents; role 1 sees documents 1 and 2, role 2 sees document 3; user 7 has role 1 */
    const float x[6 * 4] = {0, 0, 0, 0, 1, 0, 0, 0, 0, 2, 0, 0, 0, 0, 3, 0, 4, 0, 0, 0, 0, 0, 0, 5};
    const uint64_t tid[6] = {101, 102, 103, 104, 105, 106};
    const uint32_t doc[6] = {1, 1, 2, 2, 3, 3};
    const uint32_t pa_doc[3] = {1, 2, 3}, pa_role[3] = {1, 1, 2};
    const uint32_t ur_user[1] = {7}, ur_role[1] = {1};
    if (rk_load_rows(ix, x, tid, doc, 6) < 0) return fail("rk_load_rows");
    if (rk_set_doc_roles(ix, pa_doc, pa_role, 3) < 0) return fail("rk_set_doc_roles");
    if (rk_set_user_roles(ix, ur_user, ur_role, 1) < 0) return fail("rk_set_user_roles");
    if (rk_build_role_partitions(ix) < 0) return fail("rk_build_role_partitions");
    if (rk_finalize(ix) < 0) return fail("rk_finalize");
    rk_scan_opts o;
    memset(&o, 0, sizeof o);
    o.k = 8;
    o.filter = RK_FILTER_PARTITION;
    o.layout = RK_LAYOUT_ROLE;
    const float q[4] = {0, 0, 0, 0};
    rk_scan *sc = rk_beginscan(ix);
    if (!sc) return fail("rk_beginscan");
    if (rk_rescan(sc, q, 4, 7, &o) < 0) return fail("rk_rescan");
    uint64_t t;
    uint32_t d;
    double dist;
    int rc;
    while ((rc = rk_gettuple(sc, &t, &d, &dist)) == 1) printf("%llu %u %.17g\n", (unsigned long long) t, d, dist);
    if (rc < 0) return fail("rk_gettuple");
    if (rk_rescan(sc, q, 3, 7, &o) != RK_ERR_DIM) return 1;
    printf("err %s\n", rk_last_error());
    rk_endscan(sc);
    uint64_t bt[8];
    uint32_t bd[8];
    double bx[8];
    int32_t bc = 0;
    const uint32_t user = 7;
    o.filter = RK_FILTER_BITMASK;
    if (rk_search_batch(ix, 1, q, &user, &o, bt, bd, bx, &bc) < 0) return fail("rk_search_batch");
    printf("batch %d", bc);
    for (int i = 0; i < bc; i++) printf(" %llu", (unsigned long long) bt[i]);
    printf("\n");
    rk_close(ix);
    return 0;
}
