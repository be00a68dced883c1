/*
 * minifits — a tiny stand-in for the subset of the cfitsio API that the
 * reference hotpants sources call (hotpants/main.c, functions.c, vargs.c).
 *
 * TEST INFRASTRUCTURE ONLY.  cfitsio is not installed in this image, so the
 * reference program is linked against this shim to obtain oracle/_ref/hotpants
 * (see oracle/Makefile).  It is pure I/O: no result-bearing arithmetic lives
 * here.  Not part of the shipped product; nothing under zuds_pipeline_b200/
 * links or loads it.
 *
 * Supported: 2-D images with BITPIX 8/16/32/-32/-64 (+BSCALE/BZERO), header
 * cards, and one binary-table flavour (all-double columns) for the -hki table.
 */
#ifndef MINIFITS_FITSIO_H
#define MINIFITS_FITSIO_H

#include <stdio.h>

#define FLEN_CARD 81

#define BYTE_IMG     8
#define SHORT_IMG   16
#define LONG_IMG    32
#define FLOAT_IMG  -32
#define DOUBLE_IMG -64

#define TLOGICAL 14
#define TSTRING  16
#define TINT     31
#define TFLOAT   42
#define TDOUBLE  82

#define READONLY  0
#define READWRITE 1

/* status codes (cfitsio numbering where it matters) */
#define FILE_NOT_OPENED   104
#define FILE_NOT_CREATED  105
#define WRITE_ERROR       106
#define READ_ERROR        108
#define KEY_NO_EXIST      202
#define BAD_HDU_NUM       301
#define BAD_DIMEN         320
#define HEADER_FULL       2001  /* minifits only: reserve of blank cards exhausted */
#define UNSUPPORTED_OP    2002  /* minifits only */

typedef struct minifits_file fitsfile;

int fits_open_file(fitsfile **fptr, const char *filename, int iomode, int *status);
int fits_create_file(fitsfile **fptr, const char *filename, int *status);
int fits_close_file(fitsfile *fptr, int *status);

int fits_create_img(fitsfile *fptr, int bitpix, int naxis, long *naxes, int *status);
int fits_insert_img(fitsfile *fptr, int bitpix, int naxis, long *naxes, int *status);
int fits_insert_btbl(fitsfile *fptr, long naxis2, int tfields, char **ttype, char **tform,
                     char **tunit, const char *extname, long pcount, int *status);

int fits_get_img_param(fitsfile *fptr, int maxdim, int *bitpix, int *naxis, long *naxes, int *status);
int fits_get_hdu_num(fitsfile *fptr, int *hdunum);
int fits_get_num_hdus(fitsfile *fptr, int *nhdu, int *status);
int fits_get_hdrspace(fitsfile *fptr, int *nexist, int *nmore, int *status);
int fits_movabs_hdu(fitsfile *fptr, int hdunum, int *exttype, int *status);
int fits_movrel_hdu(fitsfile *fptr, int hdumov, int *exttype, int *status);

int fits_read_subset_flt(fitsfile *fptr, long group, long naxis, long *naxes, long *fpixel,
                         long *lpixel, long *inc, float nulval, float *array, int *anynul, int *status);
int fits_read_subset_int(fitsfile *fptr, long group, long naxis, long *naxes, long *fpixel,
                         long *lpixel, long *inc, int nulval, int *array, int *anynul, int *status);
int fits_read_img_flt(fitsfile *fptr, long group, long firstelem, long nelem, float nulval,
                      float *array, int *anynul, int *status);
int fits_write_subset_flt(fitsfile *fptr, long group, long naxis, long *naxes, long *fpixel,
                          long *lpixel, float *array, int *status);
int fits_write_subset_int(fitsfile *fptr, long group, long naxis, long *naxes, long *fpixel,
                          long *lpixel, int *array, int *status);
int fits_write_col_dbl(fitsfile *fptr, int colnum, long firstrow, long firstelem, long nelem,
                       double *array, int *status);
int fits_read_col(fitsfile *fptr, int datatype, int colnum, long firstrow, long firstelem,
                  long nelem, void *nulval, void *array, int *anynul, int *status);

int fits_read_key(fitsfile *fptr, int datatype, const char *keyname, void *value, char *comm, int *status);
int fits_read_key_str(fitsfile *fptr, const char *keyname, char *value, char *comm, int *status);
int fits_read_key_lng(fitsfile *fptr, const char *keyname, long *value, char *comm, int *status);
int fits_read_key_log(fitsfile *fptr, const char *keyname, int *value, char *comm, int *status);
int fits_read_record(fitsfile *fptr, int nrec, char *card, int *status);

int fits_write_record(fitsfile *fptr, const char *card, int *status);
int fits_write_key_str(fitsfile *fptr, const char *keyname, const char *value, const char *comm, int *status);
int fits_write_key_longstr(fitsfile *fptr, const char *keyname, const char *value, const char *comm, int *status);
int fits_write_key_lng(fitsfile *fptr, const char *keyname, long value, const char *comm, int *status);
int fits_write_key_flt(fitsfile *fptr, const char *keyname, float value, int decim, const char *comm, int *status);
int fits_write_key_log(fitsfile *fptr, const char *keyname, int value, const char *comm, int *status);
int fits_update_key(fitsfile *fptr, int datatype, const char *keyname, void *value, const char *comm, int *status);
int fits_update_key_flt(fitsfile *fptr, const char *keyname, float value, int decim, const char *comm, int *status);

int fits_set_bscale(fitsfile *fptr, double scale, double zero, int *status);
void fits_report_error(FILE *stream, int status);

#endif
