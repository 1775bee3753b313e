// krrt_real.cuh — the arithmetic type of the kernels that exist in two precisions (krrt_math.cuh, krrt_warp.cuh,
// krrt_batch.cuh, krrt_batch2.cuh).
//
// Those files are written in `double`.  The fp64 library compiles them as they stand.  The fp32 build (krrt_f32.cu,
// KRRT_F32 contexts: include/krrt.h `precision`) compiles THE SAME SOURCE with every `double` token replaced by `float`
// (`#define double float`, likewise double2 / double4 / make_double2 / make_double4, namespace krrt -> krrt_f32) and every
// floating literal given an `f` suffix through KR(): the fp32 planner is, operation for operation, the fp64 planner in
// binary32 — exactly what `gcc -fsingle-precision-constant` with the same substitution makes of the CPU checker, so
// the two compare bit for bit.  What crosses the C ABI stays binary64 in both builds: krrt_io_real (inputs are rounded
// to the arithmetic type when a kernel loads them, results widen exactly when it stores them).
#pragma once
typedef double krrt_io_real;  // (seen before any `#define double float`: this header has an include guard)
#ifndef KR
#define KR(x) x
#endif
