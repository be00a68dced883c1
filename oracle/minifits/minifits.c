/*
 * minifits.c — see fitsio.h.  TEST INFRASTRUCTURE ONLY (oracle/_ref build).
 *
 * Model: headers live in memory (80-byte cards, END implicit); pixel data stay
 * on disk and are moved with pread/pwrite a row at a time, like cfitsio does.
 * Files created here get RESERVE_BLOCKS of blank cards before END so that
 * keywords can be appended after the data have been written without moving
 * the data (cfitsio would shift the data instead).
 */
#define _XOPEN_SOURCE 700
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#include <fcntl.h>
#include <math.h>
#include <stdint.h>
#include <sys/types.h>
#include <sys/stat.h>
#include "fitsio.h"

#define BLK 2880
#define CARDS_PER_BLK 36
#define RESERVE_BLOCKS 40
#define MAXHDU 8

typedef struct {
    char  *cards;            /* ncards * 80 bytes, no END card */
    int    ncards, cap;
    int    hdr_blocks;       /* 0 until the HDU has a place on disk */
    off_t  hdr_off, data_off;
    long   data_bytes;
    int    dirty;
    int    is_table;
    int    bitpix, naxis;
    long   naxes[4];
    long   rowlen, nrows;
    int    tfields;
    int    have_scale;       /* fits_set_bscale override */
    double oscale, ozero;
} hdu_t;

struct minifits_file {
    int   fd, rw, nhdu, cur, placed;
    hdu_t h[MAXHDU];
};

/* ---------------------------------------------------------------- cards */

static void card_blank(char *c) { memset(c, ' ', 80); }

static void hdu_reserve(hdu_t *h, int n) {
    if (h->ncards + n > h->cap) {
        h->cap = (h->ncards + n) * 2 + 64;
        h->cards = (char *)realloc(h->cards, (size_t)h->cap * 80);
    }
}

static void hdu_append(hdu_t *h, const char *card80) {
    hdu_reserve(h, 1);
    memcpy(h->cards + (size_t)h->ncards * 80, card80, 80);
    h->ncards++;
    h->dirty = 1;
}

static int hdu_find(const hdu_t *h, const char *key) {
    char k[9];
    int i, n = (int)strlen(key);
    if (n > 8) n = 8;
    memset(k, ' ', 8); k[8] = 0;
    for (i = 0; i < n; i++) k[i] = key[i];
    for (i = 0; i < h->ncards; i++)
        if (memcmp(h->cards + (size_t)i * 80, k, 8) == 0) return i;
    return -1;
}

/* "KEYWORD = <value right/left justified> / comment" */
static void card_make(char *out, const char *key, const char *val, int is_string, const char *comm) {
    char buf[512];
    int n;
    if (is_string) {
        char q[160]; int j = 0; const char *p;
        q[j++] = '\'';
        for (p = val; *p && j < 140; p++) { q[j++] = *p; if (*p == '\'') q[j++] = '\''; }
        while (j < 9) q[j++] = ' ';
        q[j++] = '\''; q[j] = 0;
        n = snprintf(buf, sizeof buf, "%-8.8s= %-20s", key, q);
    } else {
        n = snprintf(buf, sizeof buf, "%-8.8s= %20s", key, val);
    }
    if (comm && comm[0] && n < 77) snprintf(buf + n, sizeof buf - (size_t)n, " / %s", comm);
    card_blank(out);
    n = (int)strlen(buf); if (n > 80) n = 80;
    memcpy(out, buf, (size_t)n);
}

static void hdu_put(hdu_t *h, const char *key, const char *val, int is_string, const char *comm, int update) {
    char c[80];
    int i = update ? hdu_find(h, key) : -1;
    if (i >= 0 && (!comm || !comm[0])) {
        /* keep the existing comment on update, as cfitsio does */
        char old[81]; char *s;
        memcpy(old, h->cards + (size_t)i * 80, 80); old[80] = 0;
        s = strstr(old + 10, " / ");
        if (s) { int e = 79; while (e > 0 && old[e] == ' ') old[e--] = 0; comm = s + 3; }
        card_make(c, key, val, is_string, comm);
    } else
        card_make(c, key, val, is_string, comm);
    if (i >= 0) { memcpy(h->cards + (size_t)i * 80, c, 80); h->dirty = 1; }
    else hdu_append(h, c);
}

/* parse the value field of card i into buf (strings unquoted, trailing blanks cut) */
static int hdu_value(const hdu_t *h, int i, char *buf, int buflen) {
    char c[81]; int j, k = 0;
    memcpy(c, h->cards + (size_t)i * 80, 80); c[80] = 0;
    j = 10;
    while (j < 80 && c[j] == ' ') j++;
    if (j < 80 && c[j] == '\'') {
        j++;
        while (j < 80 && k < buflen - 1) {
            if (c[j] == '\'') { if (j + 1 < 80 && c[j + 1] == '\'') { buf[k++] = '\''; j += 2; continue; } break; }
            buf[k++] = c[j++];
        }
        while (k > 0 && buf[k - 1] == ' ') k--;
    } else {
        while (j < 80 && c[j] != ' ' && c[j] != '/' && k < buflen - 1) buf[k++] = c[j++];
    }
    buf[k] = 0;
    return k;
}

static double hdu_num(const hdu_t *h, const char *key, double dflt) {
    char v[96]; int i = hdu_find(h, key);
    if (i < 0) return dflt;
    hdu_value(h, i, v, sizeof v);
    { char *p; for (p = v; *p; p++) if (*p == 'D' || *p == 'd') *p = 'E'; }
    return atof(v);
}

/* --------------------------------------------------------------- layout */

static long pad_blk(long n) { return ((n + BLK - 1) / BLK) * BLK; }

static int bytes_per_pix(int bitpix) { return (bitpix < 0 ? -bitpix : bitpix) / 8; }

static void hdu_geometry(hdu_t *h) {
    int i;
    h->is_table = 0;
    { int x = hdu_find(h, "XTENSION"); char v[32];
      if (x >= 0) { hdu_value(h, x, v, sizeof v); if (strncmp(v, "BINTABLE", 8) == 0) h->is_table = 1; } }
    h->bitpix = (int)hdu_num(h, "BITPIX", 8);
    h->naxis  = (int)hdu_num(h, "NAXIS", 0);
    for (i = 0; i < 4; i++) h->naxes[i] = 1;
    for (i = 0; i < h->naxis && i < 4; i++) {
        char k[16]; snprintf(k, sizeof k, "NAXIS%d", i + 1);
        h->naxes[i] = (long)hdu_num(h, k, 0);
    }
    if (h->naxis == 0) h->data_bytes = 0;
    else {
        long n = 1;
        for (i = 0; i < h->naxis && i < 4; i++) n *= h->naxes[i];
        h->data_bytes = n * bytes_per_pix(h->bitpix);
    }
    if (h->is_table) {
        h->rowlen  = h->naxes[0];
        h->nrows   = h->naxes[1];
        h->tfields = (int)hdu_num(h, "TFIELDS", 0);
    }
}

/* give every HDU of a freshly created file its place on disk */
static int file_place(fitsfile *f) {
    off_t off = 0; int i;
    if (f->placed) return 0;
    for (i = 0; i < f->nhdu; i++) {
        hdu_t *h = &f->h[i];
        hdu_geometry(h);
        h->hdr_blocks = (h->ncards + 1 + CARDS_PER_BLK - 1) / CARDS_PER_BLK + RESERVE_BLOCKS;
        h->hdr_off  = off;
        h->data_off = off + (off_t)h->hdr_blocks * BLK;
        off = h->data_off + pad_blk(h->data_bytes);
        h->dirty = 1;
    }
    if (ftruncate(f->fd, off) != 0) return WRITE_ERROR;
    f->placed = 1;
    return 0;
}

static int hdu_flush(fitsfile *f, hdu_t *h) {
    long total = (long)h->hdr_blocks * BLK;
    char *buf;
    if (!h->dirty) return 0;
    if ((long)(h->ncards + 1) * 80 > total) return HEADER_FULL;
    buf = (char *)malloc((size_t)total);
    memset(buf, ' ', (size_t)total);
    memcpy(buf, h->cards, (size_t)h->ncards * 80);
    memcpy(buf + total - 80, "END", 3);           /* blank reserve cards sit before END */
    if (pwrite(f->fd, buf, (size_t)total, h->hdr_off) != total) { free(buf); return WRITE_ERROR; }
    free(buf);
    h->dirty = 0;
    return 0;
}

/* ------------------------------------------------------------ open/close */

int fits_open_file(fitsfile **fptr, const char *filename, int iomode, int *status) {
    fitsfile *f; off_t off = 0; struct stat st;
    if (*status) return *status;
    f = (fitsfile *)calloc(1, sizeof *f);
    f->rw = iomode;
    f->fd = open(filename, iomode ? O_RDWR : O_RDONLY);
    if (f->fd < 0 || fstat(f->fd, &st) != 0) { free(f); return (*status = FILE_NOT_OPENED); }
    while (off < st.st_size && f->nhdu < MAXHDU) {
        hdu_t *h = &f->h[f->nhdu];
        char blk[BLK]; int done = 0, nb = 0;
        h->hdr_off = off;
        while (!done) {
            int c;
            if (pread(f->fd, blk, BLK, off + (off_t)nb * BLK) != BLK) { done = -1; break; }
            nb++;
            for (c = 0; c < CARDS_PER_BLK; c++) {
                const char *cd = blk + c * 80;
                if (memcmp(cd, "END     ", 8) == 0) { done = 1; break; }
                hdu_append(h, cd);
            }
        }
        if (done < 0) { if (f->nhdu == 0) { close(f->fd); free(f); return (*status = READ_ERROR); } h->ncards = 0; break; }
        /* drop the blank reserve in front of END */
        while (h->ncards > 0) {
            const char *cd = h->cards + (size_t)(h->ncards - 1) * 80; int k, blank = 1;
            for (k = 0; k < 80; k++) if (cd[k] != ' ') { blank = 0; break; }
            if (!blank) break;
            h->ncards--;
        }
        h->dirty = 0;
        h->hdr_blocks = nb;
        h->data_off = off + (off_t)nb * BLK;
        hdu_geometry(h);
        off = h->data_off + pad_blk(h->data_bytes);
        f->nhdu++;
    }
    f->placed = 1;
    f->cur = 0;
    *fptr = f;
    return 0;
}

int fits_create_file(fitsfile **fptr, const char *filename, int *status) {
    fitsfile *f;
    if (*status) return *status;
    f = (fitsfile *)calloc(1, sizeof *f);
    f->rw = 1;
    if (filename[0] == '!') { filename++; unlink(filename); }
    f->fd = open(filename, O_RDWR | O_CREAT | O_EXCL, 0644);
    if (f->fd < 0) { free(f); return (*status = FILE_NOT_CREATED); }
    *fptr = f;
    return 0;
}

int fits_close_file(fitsfile *f, int *status) {
    int i, st = 0;
    if (!f) return *status;
    if (f->rw) {
        st = file_place(f);
        for (i = 0; i < f->nhdu && !st; i++) st = hdu_flush(f, &f->h[i]);
    }
    for (i = 0; i < f->nhdu; i++) free(f->h[i].cards);
    close(f->fd);
    free(f);
    if (st && !*status) *status = st;
    return *status;
}

/* ------------------------------------------------------------ HDU set-up */

static void put_int(hdu_t *h, const char *key, long v, const char *comm) {
    char b[32]; snprintf(b, sizeof b, "%ld", v); hdu_put(h, key, b, 0, comm, 0);
}

static int new_img(fitsfile *f, int bitpix, int naxis, long *naxes, int *status) {
    hdu_t *h; int i; char c[81];
    if (*status) return *status;
    if (f->placed || f->nhdu >= MAXHDU) return (*status = UNSUPPORTED_OP);
    h = &f->h[f->nhdu];
    memset(h, 0, sizeof *h);
    if (f->nhdu == 0) hdu_put(h, "SIMPLE", "T", 0, "file does conform to FITS standard", 0);
    else              hdu_put(h, "XTENSION", "IMAGE", 1, "IMAGE extension", 0);
    put_int(h, "BITPIX", bitpix, "number of bits per data pixel");
    put_int(h, "NAXIS", naxis, "number of data axes");
    for (i = 0; i < naxis; i++) { char k[16]; snprintf(k, sizeof k, "NAXIS%d", i + 1); put_int(h, k, naxes[i], "length of data axis"); }
    if (f->nhdu == 0) {
        hdu_put(h, "EXTEND", "T", 0, "FITS dataset may contain extensions", 0);
        snprintf(c, sizeof c, "%-80s", "COMMENT   FITS (Flexible Image Transport System) format is defined in 'Astronomy");
        hdu_append(h, c);
        snprintf(c, sizeof c, "%-80s", "COMMENT   and Astrophysics', volume 376, page 359; bibcode: 2001A&A...376..359H");
        hdu_append(h, c);
    } else {
        put_int(h, "PCOUNT", 0, "required keyword; must = 0");
        put_int(h, "GCOUNT", 1, "required keyword; must = 1");
    }
    hdu_geometry(h);
    f->cur = f->nhdu++;
    return 0;
}

int fits_create_img(fitsfile *f, int bitpix, int naxis, long *naxes, int *status) { return new_img(f, bitpix, naxis, naxes, status); }
int fits_insert_img(fitsfile *f, int bitpix, int naxis, long *naxes, int *status) { return new_img(f, bitpix, naxis, naxes, status); }

int fits_insert_btbl(fitsfile *f, long naxis2, int tfields, char **ttype, char **tform,
                     char **tunit, const char *extname, long pcount, int *status) {
    hdu_t *h; int i;
    (void)tunit; (void)pcount;
    if (*status) return *status;
    if (f->placed || f->nhdu >= MAXHDU) return (*status = UNSUPPORTED_OP);
    if (f->nhdu == 0) {              /* a table needs an (empty) primary array in front */
        hdu_t *p = &f->h[0];
        memset(p, 0, sizeof *p);
        hdu_put(p, "SIMPLE", "T", 0, "file does conform to FITS standard", 0);
        put_int(p, "BITPIX", 8, ""); put_int(p, "NAXIS", 0, "");
        hdu_put(p, "EXTEND", "T", 0, "", 0);
        hdu_geometry(p);
        f->nhdu = 1;
    }
    h = &f->h[f->nhdu];
    memset(h, 0, sizeof *h);
    hdu_put(h, "XTENSION", "BINTABLE", 1, "binary table extension", 0);
    put_int(h, "BITPIX", 8, "8-bit bytes");
    put_int(h, "NAXIS", 2, "2-dimensional binary table");
    put_int(h, "NAXIS1", 8L * tfields, "width of table in bytes");
    put_int(h, "NAXIS2", naxis2, "number of rows in table");
    put_int(h, "PCOUNT", 0, "size of special data area");
    put_int(h, "GCOUNT", 1, "one data group");
    put_int(h, "TFIELDS", tfields, "number of fields in each row");
    for (i = 0; i < tfields; i++) {
        char k[16];
        snprintf(k, sizeof k, "TTYPE%d", i + 1); hdu_put(h, k, ttype[i], 1, "label for field", 0);
        snprintf(k, sizeof k, "TFORM%d", i + 1); hdu_put(h, k, tform[i], 1, "data format of field: 8-byte DOUBLE", 0);
    }
    if (extname && extname[0]) hdu_put(h, "EXTNAME", extname, 1, "name of this binary table extension", 0);
    hdu_geometry(h);
    f->cur = f->nhdu++;
    return 0;
}

int fits_get_img_param(fitsfile *f, int maxdim, int *bitpix, int *naxis, long *naxes, int *status) {
    hdu_t *h = &f->h[f->cur]; int i;
    if (*status) return *status;
    *bitpix = h->bitpix; *naxis = h->naxis;
    for (i = 0; i < h->naxis && i < maxdim; i++) naxes[i] = h->naxes[i];
    return 0;
}

int fits_get_hdu_num(fitsfile *f, int *hdunum) { *hdunum = f->cur + 1; return *hdunum; }
int fits_get_num_hdus(fitsfile *f, int *nhdu, int *status) { *nhdu = f->nhdu; return *status; }

int fits_get_hdrspace(fitsfile *f, int *nexist, int *nmore, int *status) {
    if (nexist) *nexist = f->h[f->cur].ncards;
    if (nmore)  *nmore  = -1;
    return *status;
}

int fits_movabs_hdu(fitsfile *f, int hdunum, int *exttype, int *status) {
    if (*status) return *status;
    if (hdunum < 1 || hdunum > f->nhdu) return (*status = BAD_HDU_NUM);
    f->cur = hdunum - 1;
    f->h[f->cur].have_scale = 0;
    if (exttype) *exttype = f->h[f->cur].is_table ? 2 : 0;
    return 0;
}

int fits_movrel_hdu(fitsfile *f, int hdumov, int *exttype, int *status) {
    return fits_movabs_hdu(f, f->cur + 1 + hdumov, exttype, status);
}

int fits_set_bscale(fitsfile *f, double scale, double zero, int *status) {
    hdu_t *h = &f->h[f->cur];
    if (*status) return *status;
    h->have_scale = 1; h->oscale = scale; h->ozero = zero;
    return 0;
}

/* -------------------------------------------------------------- pixel I/O */

static void scaling(const hdu_t *h, double *bscale, double *bzero) {
    if (h->have_scale) { *bscale = h->oscale; *bzero = h->ozero; }
    else { *bscale = hdu_num(h, "BSCALE", 1.0); *bzero = hdu_num(h, "BZERO", 0.0); }
}

static inline uint16_t sw16(uint16_t v) { return (uint16_t)((v >> 8) | (v << 8)); }
static inline uint32_t sw32(uint32_t v) { return __builtin_bswap32(v); }
static inline uint64_t sw64(uint64_t v) { return __builtin_bswap64(v); }

/* decode n raw big-endian pixels into doubles */
static void decode_row(const unsigned char *raw, int bitpix, long n, double *out) {
    long i;
    switch (bitpix) {
    case 8:   for (i = 0; i < n; i++) out[i] = raw[i]; break;
    case 16:  for (i = 0; i < n; i++) { uint16_t v; memcpy(&v, raw + 2 * i, 2); out[i] = (int16_t)sw16(v); } break;
    case 32:  for (i = 0; i < n; i++) { uint32_t v; memcpy(&v, raw + 4 * i, 4); out[i] = (int32_t)sw32(v); } break;
    case -32: for (i = 0; i < n; i++) { uint32_t v; float fv; memcpy(&v, raw + 4 * i, 4); v = sw32(v); memcpy(&fv, &v, 4); out[i] = fv; } break;
    case -64: for (i = 0; i < n; i++) { uint64_t v; double dv; memcpy(&v, raw + 8 * i, 8); v = sw64(v); memcpy(&dv, &v, 8); out[i] = dv; } break;
    default:  for (i = 0; i < n; i++) out[i] = 0;
    }
}

static int read_rect(fitsfile *f, long *fp, long *lp, float *fout, int *iout, int *status) {
    hdu_t *h = &f->h[f->cur];
    long nx = lp[0] - fp[0] + 1, ny = lp[1] - fp[1] + 1, y, i;
    int bpp = bytes_per_pix(h->bitpix);
    double bs, bz, *row;
    unsigned char *raw;
    if (*status) return *status;
    if (h->is_table || h->naxis < 2 || fp[0] < 1 || fp[1] < 1 || lp[0] > h->naxes[0] || lp[1] > h->naxes[1])
        return (*status = BAD_DIMEN);
    scaling(h, &bs, &bz);
    raw = (unsigned char *)malloc((size_t)nx * (size_t)bpp);
    row = (double *)malloc((size_t)nx * sizeof(double));
    for (y = 0; y < ny; y++) {
        off_t off = h->data_off + ((off_t)(fp[1] - 1 + y) * h->naxes[0] + (fp[0] - 1)) * bpp;
        if (pread(f->fd, raw, (size_t)nx * (size_t)bpp, off) != (ssize_t)(nx * bpp)) { *status = READ_ERROR; break; }
        if (h->bitpix == -32 && bs == 1.0 && bz == 0.0 && fout) {      /* common fast path */
            for (i = 0; i < nx; i++) { uint32_t v; memcpy(&v, raw + 4 * i, 4); v = sw32(v); memcpy(&fout[y * nx + i], &v, 4); }
            continue;
        }
        decode_row(raw, h->bitpix, nx, row);
        if (fout) for (i = 0; i < nx; i++) fout[y * nx + i] = (float)(row[i] * bs + bz);
        else      for (i = 0; i < nx; i++) { double v = row[i] * bs + bz; iout[y * nx + i] = (int)(v < 0 ? v - 0.5 : v + 0.5); }
    }
    free(raw); free(row);
    return *status;
}

int fits_read_subset_flt(fitsfile *f, long group, long naxis, long *naxes, long *fpixel, long *lpixel,
                         long *inc, float nulval, float *array, int *anynul, int *status) {
    (void)group; (void)naxis; (void)naxes; (void)inc; (void)nulval;
    if (anynul) *anynul = 0;
    return read_rect(f, fpixel, lpixel, array, NULL, status);
}

int fits_read_subset_int(fitsfile *f, long group, long naxis, long *naxes, long *fpixel, long *lpixel,
                         long *inc, int nulval, int *array, int *anynul, int *status) {
    (void)group; (void)naxis; (void)naxes; (void)inc; (void)nulval;
    if (anynul) *anynul = 0;
    return read_rect(f, fpixel, lpixel, NULL, array, status);
}

int fits_read_img_flt(fitsfile *f, long group, long firstelem, long nelem, float nulval,
                      float *array, int *anynul, int *status) {
    hdu_t *h = &f->h[f->cur]; long fp[2], lp[2];
    (void)group; (void)nulval;
    if (anynul) *anynul = 0;
    if (firstelem != 1 || nelem != h->naxes[0] * h->naxes[1]) return (*status = UNSUPPORTED_OP);
    fp[0] = fp[1] = 1; lp[0] = h->naxes[0]; lp[1] = h->naxes[1];
    return read_rect(f, fp, lp, array, NULL, status);
}

static int write_rect(fitsfile *f, long *fp, long *lp, const float *fin, const int *iin, int *status) {
    hdu_t *h = &f->h[f->cur];
    long nx = lp[0] - fp[0] + 1, ny = lp[1] - fp[1] + 1, y, i;
    int bpp, st;
    double bs, bz;
    unsigned char *raw;
    if (*status) return *status;
    if ((st = file_place(f)) != 0) return (*status = st);
    bpp = bytes_per_pix(h->bitpix);
    if (h->is_table || h->naxis < 2 || fp[0] < 1 || fp[1] < 1 || lp[0] > h->naxes[0] || lp[1] > h->naxes[1])
        return (*status = BAD_DIMEN);
    scaling(h, &bs, &bz);
    raw = (unsigned char *)malloc((size_t)nx * (size_t)bpp);
    for (y = 0; y < ny; y++) {
        off_t off = h->data_off + ((off_t)(fp[1] - 1 + y) * h->naxes[0] + (fp[0] - 1)) * bpp;
        for (i = 0; i < nx; i++) {
            double v = fin ? (double)fin[y * nx + i] : (double)iin[y * nx + i];
            if (bs != 1.0 || bz != 0.0) v = (v - bz) / bs;
            switch (h->bitpix) {
            case -32: { float fv = (float)v; uint32_t u; memcpy(&u, &fv, 4); u = sw32(u); memcpy(raw + 4 * i, &u, 4); } break;
            case -64: { uint64_t u; memcpy(&u, &v, 8); u = sw64(u); memcpy(raw + 8 * i, &u, 8); } break;
            case 32:  { double r = v < 0 ? v - 0.5 : v + 0.5; int32_t q = r > 2147483647. ? 2147483647 : (r < -2147483648. ? (int32_t)(-2147483647 - 1) : (int32_t)r);
                        uint32_t u = sw32((uint32_t)q); memcpy(raw + 4 * i, &u, 4); } break;
            case 16:  { double r = v < 0 ? v - 0.5 : v + 0.5; int16_t q = r > 32767. ? 32767 : (r < -32768. ? -32768 : (int16_t)r);
                        uint16_t u = sw16((uint16_t)q); memcpy(raw + 2 * i, &u, 2); } break;
            case 8:   { double r = v + 0.5; raw[i] = (unsigned char)(r < 0 ? 0 : (r > 255 ? 255 : r)); } break;
            }
        }
        if (pwrite(f->fd, raw, (size_t)nx * (size_t)bpp, off) != (ssize_t)(nx * bpp)) { *status = WRITE_ERROR; break; }
    }
    free(raw);
    return *status;
}

int fits_write_subset_flt(fitsfile *f, long group, long naxis, long *naxes, long *fpixel, long *lpixel,
                          float *array, int *status) {
    (void)group; (void)naxis; (void)naxes;
    return write_rect(f, fpixel, lpixel, array, NULL, status);
}

int fits_write_subset_int(fitsfile *f, long group, long naxis, long *naxes, long *fpixel, long *lpixel,
                          int *array, int *status) {
    (void)group; (void)naxis; (void)naxes;
    return write_rect(f, fpixel, lpixel, NULL, array, status);
}

int fits_write_col_dbl(fitsfile *f, int colnum, long firstrow, long firstelem, long nelem,
                       double *array, int *status) {
    hdu_t *h = &f->h[f->cur]; long r; int st;
    (void)firstelem;
    if (*status) return *status;
    if (!h->is_table || colnum < 1 || colnum > h->tfields) return (*status = BAD_DIMEN);
    if ((st = file_place(f)) != 0) return (*status = st);
    for (r = 0; r < nelem; r++) {
        uint64_t u; off_t off = h->data_off + (off_t)(firstrow - 1 + r) * h->rowlen + 8L * (colnum - 1);
        memcpy(&u, &array[r], 8); u = sw64(u);
        if (pwrite(f->fd, &u, 8, off) != 8) return (*status = WRITE_ERROR);
    }
    return 0;
}

int fits_read_col(fitsfile *f, int datatype, int colnum, long firstrow, long firstelem, long nelem,
                  void *nulval, void *array, int *anynul, int *status) {
    hdu_t *h = &f->h[f->cur]; long r; double *out = (double *)array;
    (void)firstelem; (void)nulval;
    if (anynul) *anynul = 0;
    if (*status) return *status;
    if (!h->is_table || datatype != TDOUBLE || colnum < 1 || colnum > h->tfields) return (*status = UNSUPPORTED_OP);
    for (r = 0; r < nelem; r++) {
        uint64_t u; off_t off = h->data_off + (off_t)(firstrow - 1 + r) * h->rowlen + 8L * (colnum - 1);
        if (pread(f->fd, &u, 8, off) != 8) return (*status = READ_ERROR);
        u = sw64(u); memcpy(&out[r], &u, 8);
    }
    return 0;
}

/* ------------------------------------------------------------- keywords */

int fits_read_record(fitsfile *f, int nrec, char *card, int *status) {
    hdu_t *h = &f->h[f->cur]; int e;
    if (*status) return *status;
    if (nrec < 1 || nrec > h->ncards + 1) return (*status = KEY_NO_EXIST);
    if (nrec == h->ncards + 1) { snprintf(card, FLEN_CARD, "END"); return 0; }
    memcpy(card, h->cards + (size_t)(nrec - 1) * 80, 80);
    card[80] = 0;
    for (e = 79; e >= 0 && card[e] == ' '; e--) card[e] = 0;
    return 0;
}

int fits_write_record(fitsfile *f, const char *card, int *status) {
    char c[81];
    if (*status) return *status;
    snprintf(c, sizeof c, "%-80.80s", card);
    hdu_append(&f->h[f->cur], c);
    return 0;
}

static void get_comment(const hdu_t *h, int i, char *comm) {
    char c[81]; char *s; int e;
    if (!comm) return;
    memcpy(c, h->cards + (size_t)i * 80, 80); c[80] = 0;
    for (e = 79; e >= 0 && c[e] == ' '; e--) c[e] = 0;
    s = strstr(c + 10, " / ");
    strcpy(comm, s ? s + 3 : "");
}

int fits_read_key_str(fitsfile *f, const char *keyname, char *value, char *comm, int *status) {
    hdu_t *h = &f->h[f->cur]; int i;
    if (*status) return *status;
    if ((i = hdu_find(h, keyname)) < 0) return (*status = KEY_NO_EXIST);
    hdu_value(h, i, value, 71);
    get_comment(h, i, comm);
    return 0;
}

int fits_read_key_lng(fitsfile *f, const char *keyname, long *value, char *comm, int *status) {
    hdu_t *h = &f->h[f->cur]; int i;
    if (*status) return *status;
    if ((i = hdu_find(h, keyname)) < 0) return (*status = KEY_NO_EXIST);
    *value = (long)hdu_num(h, keyname, 0);
    get_comment(h, i, comm);
    return 0;
}

int fits_read_key_log(fitsfile *f, const char *keyname, int *value, char *comm, int *status) {
    hdu_t *h = &f->h[f->cur]; int i; char v[80];
    if (*status) return *status;
    if ((i = hdu_find(h, keyname)) < 0) return (*status = KEY_NO_EXIST);
    hdu_value(h, i, v, sizeof v);
    *value = (v[0] == 'T');
    get_comment(h, i, comm);
    return 0;
}

int fits_read_key(fitsfile *f, int datatype, const char *keyname, void *value, char *comm, int *status) {
    hdu_t *h = &f->h[f->cur]; int i;
    if (*status) return *status;
    if ((i = hdu_find(h, keyname)) < 0) return (*status = KEY_NO_EXIST);
    switch (datatype) {
    case TSTRING:  hdu_value(h, i, (char *)value, 71); break;
    case TINT:     *(int *)value    = (int)hdu_num(h, keyname, 0); break;
    case TFLOAT:   *(float *)value  = (float)hdu_num(h, keyname, 0); break;
    case TDOUBLE:  *(double *)value = hdu_num(h, keyname, 0); break;
    case TLOGICAL: { char v[80]; hdu_value(h, i, v, sizeof v); *(int *)value = (v[0] == 'T'); } break;
    default: return (*status = UNSUPPORTED_OP);
    }
    get_comment(h, i, comm);
    return 0;
}

int fits_write_key_str(fitsfile *f, const char *keyname, const char *value, const char *comm, int *status) {
    if (*status) return *status;
    hdu_put(&f->h[f->cur], keyname, value, 1, comm, 0);
    return 0;
}

/* long strings use the CONTINUE convention, 67 characters a piece */
int fits_write_key_longstr(fitsfile *f, const char *keyname, const char *value, const char *comm, int *status) {
    hdu_t *h = &f->h[f->cur];
    size_t len = strlen(value), pos = 0; int first = 1;
    if (*status) return *status;
    if (len <= 68) { hdu_put(h, keyname, value, 1, comm, 0); return 0; }
    while (pos < len || first) {
        char piece[80], card[192], q[96];
        size_t n = len - pos; int more;
        if (n > 67) n = 67;
        memcpy(piece, value + pos, n); piece[n] = 0;
        pos += n; more = pos < len;
        snprintf(q, sizeof q, "'%s%s'", piece, more ? "&" : "");
        if (first) snprintf(card, sizeof card, "%-8.8s= %s", keyname, q);
        else       snprintf(card, sizeof card, "CONTINUE  %s", q);
        if (!more && comm && comm[0] && strlen(card) + strlen(comm) + 3 <= 80) { strcat(card, " / "); strcat(card, comm); }
        { char c80[81]; snprintf(c80, sizeof c80, "%-80.80s", card); hdu_append(h, c80); }
        first = 0;
    }
    return 0;
}

int fits_write_key_lng(fitsfile *f, const char *keyname, long value, const char *comm, int *status) {
    char b[32];
    if (*status) return *status;
    snprintf(b, sizeof b, "%ld", value);
    hdu_put(&f->h[f->cur], keyname, b, 0, comm, 0);
    return 0;
}

static void fmt_flt(char *b, size_t n, float value, int decim) {
    if (decim < 0) snprintf(b, n, "%.*G", -decim, (double)value);
    else           snprintf(b, n, "%.*E", decim, (double)value);
    if (!strchr(b, '.') && !strchr(b, 'E') && !strchr(b, 'N') && !strchr(b, 'I')) strncat(b, ".", n - strlen(b) - 1);
}

int fits_write_key_flt(fitsfile *f, const char *keyname, float value, int decim, const char *comm, int *status) {
    char b[64];
    if (*status) return *status;
    fmt_flt(b, sizeof b, value, decim);
    hdu_put(&f->h[f->cur], keyname, b, 0, comm, 0);
    return 0;
}

int fits_update_key_flt(fitsfile *f, const char *keyname, float value, int decim, const char *comm, int *status) {
    char b[64];
    if (*status) return *status;
    fmt_flt(b, sizeof b, value, decim);
    hdu_put(&f->h[f->cur], keyname, b, 0, comm, 1);
    return 0;
}

int fits_write_key_log(fitsfile *f, const char *keyname, int value, const char *comm, int *status) {
    if (*status) return *status;
    hdu_put(&f->h[f->cur], keyname, value ? "T" : "F", 0, comm, 0);
    return 0;
}

int fits_update_key(fitsfile *f, int datatype, const char *keyname, void *value, const char *comm, int *status) {
    char b[64];
    if (*status) return *status;
    switch (datatype) {
    case TSTRING:  hdu_put(&f->h[f->cur], keyname, (const char *)value, 1, comm, 1); break;
    case TINT:     snprintf(b, sizeof b, "%d", *(int *)value); hdu_put(&f->h[f->cur], keyname, b, 0, comm, 1); break;
    case TLOGICAL: hdu_put(&f->h[f->cur], keyname, *(int *)value ? "T" : "F", 0, comm, 1); break;
    case TFLOAT:   fmt_flt(b, sizeof b, *(float *)value, -7); hdu_put(&f->h[f->cur], keyname, b, 0, comm, 1); break;
    case TDOUBLE:  snprintf(b, sizeof b, "%.15G", *(double *)value); hdu_put(&f->h[f->cur], keyname, b, 0, comm, 1); break;
    default: return (*status = UNSUPPORTED_OP);
    }
    return 0;
}

void fits_report_error(FILE *stream, int status) {
    if (status) fprintf(stream, "minifits: FITS I/O error, status = %d\n", status);
}
