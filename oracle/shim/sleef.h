/* oracle/shim/sleef.h — TEST INFRASTRUCTURE.
 * Prototypes of the SLEEF 3.6.0 AVX2 entry points that the reference's
 * cl-waffe2-simd/kernels/simd/x86_64/avx2.h:264-332 calls.  Stands in for the header SLEEF's
 * cmake build would generate; the definitions come from SLEEF's own sleefsimd{sp,dp}.c. */
#pragma once
#include <immintrin.h>
__m256  Sleef_sinf8_u10avx2(__m256);
__m256d Sleef_sind4_u10avx2(__m256d);
__m256  Sleef_cosf8_u10avx2(__m256);
__m256d Sleef_cosd4_u10avx2(__m256d);
__m256  Sleef_tanf8_u10avx2(__m256);
__m256d Sleef_tand4_u10avx2(__m256d);
__m256  Sleef_asinf8_u10avx2(__m256);
__m256d Sleef_asind4_u10avx2(__m256d);
__m256  Sleef_acosf8_u10avx2(__m256);
__m256d Sleef_acosd4_u10avx2(__m256d);
__m256  Sleef_atanf8_u10avx2(__m256);
__m256d Sleef_atand4_u10avx2(__m256d);
__m256  Sleef_sinhf8_u10avx2(__m256);
__m256d Sleef_sinhd4_u10avx2(__m256d);
__m256  Sleef_coshf8_u10avx2(__m256);
__m256d Sleef_coshd4_u10avx2(__m256d);
__m256  Sleef_tanhf8_u10avx2(__m256);
__m256d Sleef_tanhd4_u10avx2(__m256d);
__m256  Sleef_asinhf8_u10avx2(__m256);
__m256d Sleef_asinhd4_u10avx2(__m256d);
__m256  Sleef_acoshf8_u10avx2(__m256);
__m256d Sleef_acoshd4_u10avx2(__m256d);
__m256  Sleef_atanhf8_u10avx2(__m256);
__m256d Sleef_atanhd4_u10avx2(__m256d);
__m256  Sleef_logf8_u10avx2(__m256);
__m256d Sleef_logd4_u10avx2(__m256d);
__m256  Sleef_log1pf8_u10avx2(__m256);
__m256d Sleef_log1pd4_u10avx2(__m256d);
__m256  Sleef_log10f8_u10avx2(__m256);
__m256d Sleef_log10d4_u10avx2(__m256d);
__m256  Sleef_log2f8_u10avx2(__m256);
__m256d Sleef_log2d4_u10avx2(__m256d);
__m256  Sleef_expf8_u10avx2(__m256);
__m256d Sleef_expd4_u10avx2(__m256d);
__m256  Sleef_cbrtf8_u10avx2(__m256);
__m256d Sleef_cbrtd4_u10avx2(__m256d);
__m256  Sleef_sqrtf8_u05avx2(__m256);
__m256d Sleef_sqrtd4_u05avx2(__m256d);
__m256  Sleef_powf8_u10avx2(__m256, __m256);
__m256d Sleef_powd4_u10avx2(__m256d, __m256d);
__m256  Sleef_fabsf8_avx2(__m256);
__m256d Sleef_fabsd4_avx2(__m256d);
