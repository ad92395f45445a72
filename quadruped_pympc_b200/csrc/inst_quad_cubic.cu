// four-lanes-per-sample tick kernel, parametrisation CUBIC
#include "qpmpc_inst.cuh"
#include "qpmpc_quad.cuh"
extern "C" void qp_inst_quad_cubic(qp::KernelFn quad[2][3]) {
    QP_FILL(quad, tick_kernel_quad, qp::CUBIC)
}
