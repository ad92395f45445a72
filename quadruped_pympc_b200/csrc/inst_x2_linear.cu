// two-samples-per-thread packed-FP32 tick kernel, parametrisation LINEAR
#include "qpmpc_inst.cuh"
#include "qpmpc_x2.cuh"
extern "C" void qp_inst_x2_linear(qp::KernelFn x2[3]) {
    QP_FILL_X2(x2, qp::LINEAR)
}
