// sample-per-thread tick kernel + shard finish kernel, parametrisation LINEAR
#include "qpmpc_inst.cuh"
#include "qpmpc_tick.cuh"
#include "qpmpc_tail.cuh"
extern "C" void qp_inst_tick_linear(qp::KernelFn tick[2][3], qp::KernelFn fin[2][3], qp::KernelFn tick_h[2][3], qp::KernelFn tail[2][3]) {
    QP_FILL(tick, tick_kernel, qp::LINEAR)
    QP_FILL(fin, finish_kernel, qp::LINEAR)
    QP_FILL(tick_h, tick_kernel_h, qp::LINEAR)
    QP_FILL(tail, tail_kernel, qp::LINEAR)
}
