// ASan fuzz of the ivfflat page reader alone (host code): mutations of a valid relation must never touch memory outside it.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>
#include "rk_host.h"
int rk_fail(int code, const char *fmt, ...) { (void) fmt; return code; }
#include "rk_pgpage.cu"
int main(int argc, char **argv)
{
    FILE *f = fopen(argv[1], "rb");
    std::vector<unsigned char> rel;
    unsigned char buf[8192];
    size_t n;
    while ((n = fread(buf, 1, sizeof buf, f)) > 0) rel.insert(rel.end(), buf, buf + n);
    fclose(f);
    std::mt19937_64 rng(7);
    const size_t B = 8192, nblk = rel.size() / B;
    long ok = 0, bad = 0;
    for (int it = 0; it < atoi(argv[2]); it++) {
        // exact-size heap copy: any read past the end trips ASan
        unsigned char *b = (unsigned char *) malloc(rel.size());
        memcpy(b, rel.data(), rel.size());
        int muts = 1 + rng() % 4;
        for (int m = 0; m < muts; m++) {
            size_t blk = rng() % nblk;
            switch (rng() % 5) {
            case 0: { size_t off = blk * B + 12 + 2 * (rng() % 4); b[off] = rng(); b[off + 1] = rng(); break; }
            case 1: { size_t off = blk * B + 24 + 4 * (rng() % 150); for (int i = 0; i < 4; i++) b[off + i] = rng(); break; }
            case 2: { size_t off = blk * B + B - 8 + rng() % 8; b[off] = rng(); break; }
            case 3: { size_t off = blk * B + 24 + rng() % (B - 32); b[off] = rng(); break; }
            default: { size_t off = blk * B + rng() % (B - 16); for (int i = 0; i < 16; i++) b[off + i] = rng(); }
            }
        }
        rk_ivfpages *p = nullptr;
        if (rk_ivfpages_open(b, rel.size(), &p) == 0) {
            int32_t dim, lists; uint64_t tuples, unused;
            rk_ivfpages_info(p, &dim, &lists, &tuples, &unused);
            std::vector<float> c((size_t) lists * dim), x((size_t) tuples * dim);
            std::vector<uint32_t> ls((size_t) lists + 1), rl(tuples);
            std::vector<uint64_t> tid(tuples);
            rk_ivfpages_read(p, c.data(), ls.data(), rl.data(), x.data(), tid.data());
            rk_ivfpages_close(p);
            ok++;
        } else bad++;
        free(b);
    }
    printf("ok %ld rejected %ld\n", ok, bad);
    return 0;
}
