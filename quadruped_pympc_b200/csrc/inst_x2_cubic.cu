// two-samples-per-thread packed-FP32 tick kernel, parametrisation CUBIC
#include "qpmpc_inst.cuh"
#include "qpmpc_x2.cuh"
extern "C" void qp_inst_x2_cubic(qp::KernelFn x2[3]) {
    QP_FILL_X2(x2, qp::CUBIC)
}
