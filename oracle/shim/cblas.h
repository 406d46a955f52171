/* TEST INFRASTRUCTURE ONLY (oracle/): minimal CBLAS declarations so that the
 * unmodified reference sources under /root/reference/src (which
 * `#include <cblas.h>`, src/tensor.cpp:24) can be compiled in an image that has
 * no system BLAS headers. The three entry points the reference calls
 * (src/tensor.cpp:973,995,1018,1039) are mapped onto the `scipy_`-prefixed
 * LP64 symbols of the OpenBLAS build bundled with scipy
 * (site-packages/scipy.libs/libscipy_openblas-*.so). Nothing in the product
 * (qflex_b200/) includes this header. */
#pragma once
#ifdef __cplusplus
extern "C" {
#endif
typedef struct { float real, imag; } openblas_complex_float;
enum CBLAS_ORDER { CblasRowMajor = 101, CblasColMajor = 102 };
enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113, CblasConjNoTrans = 114 };
void scipy_cblas_cgemm(enum CBLAS_ORDER, enum CBLAS_TRANSPOSE, enum CBLAS_TRANSPOSE,
                       int M, int N, int K, const void* alpha, const void* A, int lda,
                       const void* B, int ldb, const void* beta, void* C, int ldc);
void scipy_cblas_cgemv(enum CBLAS_ORDER, enum CBLAS_TRANSPOSE, int m, int n,
                       const void* alpha, const void* a, int lda, const void* x, int incx,
                       const void* beta, void* y, int incy);
void scipy_cblas_cdotu_sub(int n, const void* x, int incx, const void* y, int incy, void* ret);
void scipy_openblas_set_num_threads(int);
int scipy_openblas_get_num_threads(void);
#define cblas_cgemm scipy_cblas_cgemm
#define cblas_cgemv scipy_cblas_cgemv
#define cblas_cdotu_sub scipy_cblas_cdotu_sub
#ifdef __cplusplus
}
#endif
