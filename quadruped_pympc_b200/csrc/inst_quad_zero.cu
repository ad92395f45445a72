// four-lanes-per-sample tick kernel, parametrisation ZERO_ORDER
#include "qpmpc_inst.cuh"
#include "qpmpc_quad.cuh"
extern "C" void qp_inst_quad_zero(qp::KernelFn quad[2][3]) {
    QP_FILL(quad, tick_kernel_quad, qp::ZERO_ORDER)
}
