/*
 * pgv_ref_wrap.c — exported shims appended AFTER the reference's own distance loops.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.c).  oracle/Makefile pipes
 *     #include <math.h> ; #define VECTOR_TARGET_CLONES ;
 *     sed -n '549,563p;596,606p' /root/reference/pgvector/src/vector.c ;
 *     this file
 * into gcc on stdin, so the translation unit holds the reference's `static float
 * VectorL2SquaredDistance` and `VectorInnerProduct` verbatim, read from where they lie — no
 * reference source is copied into this repository; only oracle/_ref/*.so is produced.
 */
float pgvref_l2sq(int dim, float *a, float *b) { return VectorL2SquaredDistance(dim, a, b); }
float pgvref_ip(int dim, float *a, float *b) { return VectorInnerProduct(dim, a, b); }
/* vector.c:577 and :635 */
double pgvref_l2_distance(int dim, float *a, float *b) { return sqrt((double) VectorL2SquaredDistance(dim, a, b)); }
double pgvref_neg_ip(int dim, float *a, float *b) { return (double) -VectorInnerProduct(dim, a, b); }
