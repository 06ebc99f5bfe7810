/* make_h5_fixtures.c — writes the HDF5 fixtures of tests/golden/h5/ with libhdf5 ITSELF (the
 * reference's vendored copy, /root/reference/3rdparty/src/hdf5, built out of tree by
 * oracle/Makefile target `h5fixtures` into oracle/_ref/).  Test infrastructure: the files pin
 * gsmfretda_b200/host/h5lite.hpp — a reader written from the format specification — against what
 * the real library writes.  Nothing here is linked into the product.
 *
 *   make_h5_fixtures <outdir> [bursts.bin]
 *
 * Every fixture holds the datasets loadhdf5 reads (/root/reference/src/loadHdf5.cpp:37-71) with
 * formula-generated content (tests/test_hdf5_loader.py recomputes it) in a different on-disk
 * form; bursts.bin (GSMB, gsmfretda_b200.synth.save_bin) is converted to from_bin.h5.
 */
#include <hdf5.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define NB 37      /* bursts */
#define NP 1000    /* photons */

static int64_t g_start[NB], g_stop[NB], g_istart[NB], g_istop[NB], g_times[NP];
static unsigned char g_ad[NP], g_dd[NP];
static float g_T[NB], g_S[NB];
static const float g_clk = 1.25e-8f, g_bga = 1500.0f, g_bgd = 2500.5f;

static void fill(void)
{
    for (int i = 0; i < NP; ++i) {
        g_times[i] = 3000000000000ll + 977ll * i + (i * i) % 13;   /* > 2^41: needs all 64 bits */
        g_ad[i] = (unsigned char)((i * 7 + 3) % 5 == 0);
        g_dd[i] = (unsigned char)(!g_ad[i] && (i % 3) != 0);
    }
    for (int b = 0; b < NB; ++b) {
        g_istart[b] = 27ll * b;
        g_istop[b] = 27ll * b + 20 + (b % 7);
        g_start[b] = g_times[g_istart[b]];
        g_stop[b] = g_times[g_istop[b]];
        g_T[b] = (float)(g_stop[b] - g_start[b]) * g_clk;
        g_S[b] = (float)((b * 37) % 101) / 100.0f;
    }
}

struct opts {
    int chunk;          /* 0: contiguous */
    int deflate;        /* level, 0: none */
    int shuffle, fletcher;
    int early;          /* allocate chunks early (implicit index with the latest format) */
    int compact_scalars;
    int scalar_as_array; /* clk_p etc. as shape (1,) */
    int wide_types;     /* times int32-able? no: T/SgDivSr as float64, masks as enum, istart as int32, start big-endian */
    int unlimited;      /* maxshape unlimited (extensible array with the latest format) */
};

static hid_t dcpl_for(const struct opts* o, hsize_t n)
{
    hid_t p = H5Pcreate(H5P_DATASET_CREATE);
    if (o->chunk > 0 && n > 0) {
        hsize_t c = (hsize_t)o->chunk;
        if (c > n && !o->unlimited) c = n;   /* a fixed dimension cannot be smaller than its chunk */
        H5Pset_chunk(p, 1, &c);
        if (o->shuffle) H5Pset_shuffle(p);
        if (o->deflate) H5Pset_deflate(p, (unsigned)o->deflate);
        if (o->fletcher) H5Pset_fletcher32(p);
        if (o->early) H5Pset_alloc_time(p, H5D_ALLOC_TIME_EARLY);
    }
    return p;
}

static void put(hid_t loc, const char* name, hid_t ftype, hid_t mtype, hsize_t n, const void* data, const struct opts* o)
{
    hsize_t maxn = o->unlimited ? H5S_UNLIMITED : n;
    hid_t sp = H5Screate_simple(1, &n, &maxn);
    hid_t p = dcpl_for(o, n);
    hid_t d = H5Dcreate2(loc, name, ftype, sp, H5P_DEFAULT, p, H5P_DEFAULT);
    if (d < 0) { fprintf(stderr, "H5Dcreate2 %s failed\n", name); exit(1); }
    if (n > 0) H5Dwrite(d, mtype, H5S_ALL, H5S_ALL, H5P_DEFAULT, data);
    H5Dclose(d); H5Pclose(p); H5Sclose(sp);
}

static void put_scalar(hid_t loc, const char* name, float v, const struct opts* o)
{
    if (o->scalar_as_array) {
        struct opts plain = {0};
        put(loc, name, H5T_IEEE_F32LE, H5T_NATIVE_FLOAT, 1, &v, &plain);
        return;
    }
    hid_t sp = H5Screate(H5S_SCALAR);
    hid_t p = H5Pcreate(H5P_DATASET_CREATE);
    if (o->compact_scalars) H5Pset_layout(p, H5D_COMPACT);
    hid_t ft = o->wide_types ? H5T_IEEE_F64BE : H5T_IEEE_F32LE;
    hid_t d = H5Dcreate2(loc, name, ft, sp, H5P_DEFAULT, p, H5P_DEFAULT);
    H5Dwrite(d, H5T_NATIVE_FLOAT, H5S_ALL, H5S_ALL, H5P_DEFAULT, &v);
    H5Dclose(d); H5Pclose(p); H5Sclose(sp);
}

static hid_t bool_enum(void)   /* what h5py stores numpy.bool_ as */
{
    hid_t t = H5Tenum_create(H5T_STD_I8LE);
    signed char f = 0, tr = 1;
    H5Tenum_insert(t, "FALSE", &f);
    H5Tenum_insert(t, "TRUE", &tr);
    return t;
}

static void write_all(hid_t file, const struct opts* o, int64_t nb, int64_t np, const int64_t* start,
                      const int64_t* stop, const int64_t* istart, const int64_t* istop, const int64_t* times,
                      const unsigned char* ad, const unsigned char* dd, const float* T, const float* S, float clk,
                      float bga, float bgd, int extra_links)
{
    hid_t g = H5Gcreate2(file, "/sub_bursts_l", H5P_DEFAULT, H5P_DEFAULT, H5P_DEFAULT);
    hid_t i64 = H5T_STD_I64LE;
    put(g, "start", o->wide_types ? H5T_STD_I64BE : i64, H5T_NATIVE_INT64, (hsize_t)nb, start, o);
    put(g, "stop", i64, H5T_NATIVE_INT64, (hsize_t)nb, stop, o);
    put(g, "istart", o->wide_types ? H5T_STD_I32LE : i64, H5T_NATIVE_INT64, (hsize_t)nb, istart, o);
    put(g, "istop", o->wide_types ? H5T_STD_U16LE : i64, H5T_NATIVE_INT64, (hsize_t)nb, istop, o);
    H5Gclose(g);
    put(file, "times", i64, H5T_NATIVE_INT64, (hsize_t)np, times, o);
    if (o->wide_types) {
        hid_t be = bool_enum();
        put(file, "mask_ad", be, be, (hsize_t)np, ad, o);
        put(file, "mask_dd", be, be, (hsize_t)np, dd, o);
        H5Tclose(be);
    } else {
        put(file, "mask_ad", H5T_STD_U8LE, H5T_NATIVE_UCHAR, (hsize_t)np, ad, o);
        put(file, "mask_dd", H5T_STD_U8LE, H5T_NATIVE_UCHAR, (hsize_t)np, dd, o);
    }
    hid_t ff = o->wide_types ? H5T_IEEE_F64LE : H5T_IEEE_F32LE;
    put(file, "T_burst_duration", ff, H5T_NATIVE_FLOAT, (hsize_t)nb, T, o);
    put(file, "SgDivSr", ff, H5T_NATIVE_FLOAT, (hsize_t)nb, S, o);
    put_scalar(file, "clk_p", clk, o);
    put_scalar(file, "bg_ad_rate", bga, o);
    put_scalar(file, "bg_dd_rate", bgd, o);
    /* more links in the root group than a new-style group keeps in its header (8): forces the
       dense storage when the file uses the latest format */
    for (int i = 0; i < extra_links; ++i) {
        char nm[32];
        snprintf(nm, sizeof nm, "extra_%02d", i);
        int v = i;
        struct opts plain = {0};
        put(file, nm, H5T_STD_I32LE, H5T_NATIVE_INT, 1, &v, &plain);
    }
    /* something that is not numeric, for the error path */
    {
        hid_t st = H5Tcopy(H5T_C_S1);
        H5Tset_size(st, 8);
        hsize_t one = 1;
        hid_t sp = H5Screate_simple(1, &one, NULL);
        hid_t d = H5Dcreate2(file, "a_string", st, sp, H5P_DEFAULT, H5P_DEFAULT, H5P_DEFAULT);
        H5Dwrite(d, st, H5S_ALL, H5S_ALL, H5P_DEFAULT, "abcdefg");
        H5Dclose(d); H5Sclose(sp); H5Tclose(st);
    }
}

static void make(const char* dir, const char* name, const struct opts* o, int latest, hsize_t userblock, int extra_links)
{
    char path[1024];
    snprintf(path, sizeof path, "%s/%s", dir, name);
    hid_t fapl = H5Pcreate(H5P_FILE_ACCESS);
    if (latest) H5Pset_libver_bounds(fapl, H5F_LIBVER_LATEST, H5F_LIBVER_LATEST);
    else H5Pset_libver_bounds(fapl, H5F_LIBVER_EARLIEST, H5F_LIBVER_V18);
    hid_t fcpl = H5Pcreate(H5P_FILE_CREATE);
    if (userblock) H5Pset_userblock(fcpl, userblock);
    hid_t f = H5Fcreate(path, H5F_ACC_TRUNC, fcpl, fapl);
    if (f < 0) { fprintf(stderr, "cannot create %s\n", path); exit(1); }
    write_all(f, o, NB, NP, g_start, g_stop, g_istart, g_istop, g_times, g_ad, g_dd, g_T, g_S, g_clk, g_bga, g_bgd, extra_links);
    H5Fclose(f); H5Pclose(fcpl); H5Pclose(fapl);
    printf("wrote %s\n", path);
}

static int rd(FILE* f, void* p, size_t sz, size_t n) { return n == 0 || fread(p, sz, n, f) == n; }

static void convert_bin(const char* dir, const char* bin)
{
    FILE* f = fopen(bin, "rb");
    if (!f) { perror(bin); exit(1); }
    char magic[4]; uint32_t ver; int64_t B, N; float hdr[4];
    if (fread(magic, 1, 4, f) != 4 || memcmp(magic, "GSMB", 4) || fread(&ver, 4, 1, f) != 1 || fread(&B, 8, 1, f) != 1 ||
        fread(&N, 8, 1, f) != 1 || fread(hdr, 4, 4, f) != 4) { fprintf(stderr, "bad GSMB header\n"); exit(1); }
    int64_t *start = malloc(8 * B), *stop = malloc(8 * B), *istart = malloc(8 * B), *istop = malloc(8 * B);
    int32_t* ph = malloc(4 * B);
    float *T = malloc(4 * B), *S = malloc(4 * B);
    int64_t *ad = malloc(8 * N), *dd = malloc(8 * N), *times = malloc(8 * N);
    unsigned char *mad = malloc(N), *mdd = malloc(N);
    if (!rd(f, start, 8, B) || !rd(f, stop, 8, B) || !rd(f, istart, 8, B) || !rd(f, ph, 4, B) || !rd(f, T, 4, B) ||
        !rd(f, S, 4, B) || !rd(f, ad, 8, N) || !rd(f, dd, 8, N)) { fprintf(stderr, "short GSMB file\n"); exit(1); }
    fclose(f);
    for (int64_t b = 0; b < B; ++b) istop[b] = istart[b] + ph[b] - 1;
    for (int64_t i = 0; i < N; ++i) { mad[i] = ad[i] != 0; mdd[i] = dd[i] != 0; times[i] = ad[i] ? ad[i] : dd[i]; }
    char path[1024];
    snprintf(path, sizeof path, "%s/from_bin.h5", dir);
    struct opts o = {0};
    o.chunk = 4096; o.deflate = 4; o.shuffle = 1;
    hid_t fapl = H5Pcreate(H5P_FILE_ACCESS);
    H5Pset_libver_bounds(fapl, H5F_LIBVER_EARLIEST, H5F_LIBVER_V18);
    hid_t file = H5Fcreate(path, H5F_ACC_TRUNC, H5P_DEFAULT, fapl);
    write_all(file, &o, B, N, start, stop, istart, istop, times, mad, mdd, T, S, hdr[0], hdr[1], hdr[2], 0);
    H5Fclose(file); H5Pclose(fapl);
    printf("wrote %s (%lld bursts, %lld photons)\n", path, (long long)B, (long long)N);
}

int main(int argc, char** argv)
{
    if (argc < 2) { fprintf(stderr, "usage: %s outdir [bursts.bin]\n", argv[0]); return 2; }
    fill();
    struct opts o;
    memset(&o, 0, sizeof o);
    make(argv[1], "earliest_contiguous.h5", &o, 0, 0, 0);           /* h5py defaults */
    make(argv[1], "userblock512.h5", &o, 0, 512, 0);
    o.chunk = 96; o.deflate = 6; o.shuffle = 1; o.fletcher = 1;
    make(argv[1], "earliest_chunked_gzip.h5", &o, 0, 0, 0);          /* compression="gzip", shuffle, fletcher32 */
    memset(&o, 0, sizeof o);
    o.wide_types = 1; o.chunk = 64;
    make(argv[1], "other_types.h5", &o, 0, 0, 0);                    /* BE, int32, uint16, float64, enum bool */
    memset(&o, 0, sizeof o);
    o.compact_scalars = 1;
    make(argv[1], "latest_dense.h5", &o, 1, 0, 12);                  /* superblock 3, OHDR v2, dense root group */
    o.chunk = 128; o.deflate = 1;
    make(argv[1], "latest_fixed_array.h5", &o, 1, 0, 0);             /* filtered chunks, fixed array index */
    o.chunk = 4096;
    make(argv[1], "latest_single_chunk.h5", &o, 1, 0, 0);            /* one filtered chunk per dataset */
    memset(&o, 0, sizeof o);
    o.chunk = 100; o.early = 1; o.scalar_as_array = 1;
    make(argv[1], "latest_implicit.h5", &o, 1, 0, 0);                /* early allocation, no filter: implicit index */
    memset(&o, 0, sizeof o);
    o.chunk = 100; o.unlimited = 1;
    make(argv[1], "latest_unlimited.h5", &o, 1, 0, 0);               /* extensible array: must be refused by name */
    make(argv[1], "earliest_unlimited.h5", &o, 0, 0, 0);             /* same request, version-1 B-tree: readable */
    if (argc > 2) convert_bin(argv[1], argv[2]);
    return 0;
}
