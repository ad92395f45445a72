// four-lanes-per-sample tick kernel, parametrisation LINEAR
#include "qpmpc_inst.cuh"
#include "qpmpc_quad.cuh"
extern "C" void qp_inst_quad_linear(qp::KernelFn quad[2][3]) {
    QP_FILL(quad, tick_kernel_quad, qp::LINEAR)
}
