// sample-per-thread tick kernel + shard finish kernel, parametrisation ZERO_ORDER
#include "qpmpc_inst.cuh"
#include "qpmpc_tick.cuh"
#include "qpmpc_tail.cuh"
extern "C" void qp_inst_tick_zero(qp::KernelFn tick[2][3], qp::KernelFn fin[2][3], qp::KernelFn tick_h[2][3], qp::KernelFn tail[2][3]) {
    QP_FILL(tick, tick_kernel, qp::ZERO_ORDER)
    QP_FILL(fin, finish_kernel, qp::ZERO_ORDER)
    QP_FILL(tick_h, tick_kernel_h, qp::ZERO_ORDER)
    QP_FILL(tail, tail_kernel, qp::ZERO_ORDER)
}
