// qpmpc_inst.cuh -- explicit instantiation units.  The 3 x 2 x 3 (parametrisation x adaptive x method) variants of the
// tick kernels (sample per thread, the same with noise helper threads, four lanes per sample) are compiled in six translation units (inst_*.cu) so that the build parallelises; qpmpc_api.cu
// collects the function pointers through these accessors.
#pragma once
#include "qpmpc_kernels.cuh"

namespace qp {
typedef void (*KernelFn)(const KParams);
}

extern "C" {
void qp_inst_tick_zero(qp::KernelFn tick[2][3], qp::KernelFn fin[2][3], qp::KernelFn tick_h[2][3], qp::KernelFn tail[2][3]);
void qp_inst_tick_linear(qp::KernelFn tick[2][3], qp::KernelFn fin[2][3], qp::KernelFn tick_h[2][3], qp::KernelFn tail[2][3]);
void qp_inst_tick_cubic(qp::KernelFn tick[2][3], qp::KernelFn fin[2][3], qp::KernelFn tick_h[2][3], qp::KernelFn tail[2][3]);
void qp_inst_quad_zero(qp::KernelFn quad[2][3]);
void qp_inst_quad_linear(qp::KernelFn quad[2][3]);
void qp_inst_quad_cubic(qp::KernelFn quad[2][3]);
void qp_inst_x2_zero(qp::KernelFn x2[3]);
void qp_inst_x2_linear(qp::KernelFn x2[3]);
void qp_inst_x2_cubic(qp::KernelFn x2[3]);
}

#define QP_FILL(T, K, P)                                                                        \
    T[0][0] = qp::K<P, false, qp::RANDOM_SAMPLING>; T[0][1] = qp::K<P, false, qp::MPPI>;        \
    T[0][2] = qp::K<P, false, qp::CEM_MPPI>; T[1][0] = qp::K<P, true, qp::RANDOM_SAMPLING>;     \
    T[1][1] = qp::K<P, true, qp::MPPI>; T[1][2] = qp::K<P, true, qp::CEM_MPPI>;
#define QP_FILL_X2(T, P)                                                                       \
    T[0] = qp::tick_kernel_x2<P, qp::RANDOM_SAMPLING>; T[1] = qp::tick_kernel_x2<P, qp::MPPI>;     \
    T[2] = qp::tick_kernel_x2<P, qp::CEM_MPPI>;
