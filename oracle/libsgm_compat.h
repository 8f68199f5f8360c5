/* Force-included when oracle/Makefile compiles the reference's vendored libSGM: its kernels call the warp shuffles without
 * _sync, which the toolkit declares only below sm_70.  Full-mask _sync forms are what those calls meant (every call site
 * runs with all 32 lanes active). */
#pragma once
#define __shfl(v, l) __shfl_sync(0xffffffffu, (v), (l))
#define __shfl_xor(v, m) __shfl_xor_sync(0xffffffffu, (v), (m))
#define __shfl_up(v, d, w) __shfl_up_sync(0xffffffffu, (v), (d), (w))
#define __shfl_down(v, d, w) __shfl_down_sync(0xffffffffu, (v), (d), (w))
