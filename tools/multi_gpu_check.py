"""Multi-GPU check of the sample-sharded controller (SURVEY 8e / BASELINE configs[3]); run under torchrun, one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29533 \
        tools/multi_gpu_check.py [--ticks 300] [--out gpurun_out/sharded_G.jsonl]

For exchange in {peer (record exchange fused into the tick kernel over NVLink peer memory), nccl (partial kernel +
ncclAllGather + finish kernel)} and N in 1k..1M samples:
  * parity: the sharded tick (in-kernel Philox, global sample indices) against the unsharded tick of the same controller
    on the same inputs (rank 0 runs the unsharded one): best_index exact, best_cost / best / GRFs within the
    north_star's FP32 tolerance (merge order differs), and all ranks bit-identical among themselves;
  * latency: p50 / p99 of the full boundary call (host buffers), max over ranks of the mean.
Prints one JSON line per case on rank 0.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadruped_pympc_b200 as qp
from quadruped_pympc_b200 import synthetic
from quadruped_pympc_b200.distributed import make_sharded_controller


def tick(ctrl, sc, best, sigma=None):
    if ctrl.sampling_method == 'cem_mppi':
        return ctrl.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, ctrl.master_key,
                                           sigma, sc['phase_signal'], sc['nominal_step_freq'], 1)
    return ctrl.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, ctrl.master_key,
                                       sc['phase_signal'], sc['nominal_step_freq'], 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--ticks', type=int, default=300)
    ap.add_argument('--out', default='')
    ap.add_argument('--sizes', default='1000,10000,100000,1000000')
    ap.add_argument('--exchange', default='peer,nccl')
    ap.add_argument('--first-case', type=int, default=0, help='skip the cases before this index')
    args = ap.parse_args()
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    qp.build()
    lines = []
    cases = [('go2', 'trot', 'mppi', 'cubic_spline', False, int(n)) for n in args.sizes.split(',')]
    cases += [('go2', 'trot', 'random_sampling', 'zero_order', False, 10000),
              ('hyqreal1', 'pace', 'cem_mppi', 'cubic_spline', True, 50000),
              ('go2', 'trot', 'cem_mppi', 'linear_spline', False, 10000)]
    t_start = time.perf_counter()
    for exchange in args.exchange.split(','):
        for robot, gait, method, param, adaptive, N in cases[args.first_case:]:
            cfg = qp.SamplerConfig.for_robot(robot, sampling_method=method, control_parametrization=param,
                                             num_parallel_computations=N, optimize_step_freq=adaptive, device=local)
            ctrl = make_sharded_controller(cfg, exchange=exchange)
            sc = synthetic.scenario(robot, gait, seed=7)
            P = ctrl.num_control_parameters
            best0 = np.zeros(P, np.float32)
            sigma = np.full(P, 3.0, np.float32)
            ctrl.with_newkey()
            out = tick(ctrl, sc, best0, sigma)
            # ---- parity vs the unsharded controller (rank 0), and bit-identity across ranks ----
            ok, detail = True, {}
            if rank == 0:
                solo = qp.Sampling_MPC(cfg)
                solo.master_key = ctrl.master_key.copy()
                ref = tick(solo, sc, best0, sigma)
                detail['best_index'] = [int(ctrl.last_best_index), int(solo.last_best_index)]
                detail['best_cost_rel'] = float(abs(float(out[4]) - float(ref[4])) / abs(float(ref[4])))
                detail['best_maxabs'] = float(np.max(np.abs(np.asarray(out[3]) - np.asarray(ref[3]))))
                detail['grf_maxabs'] = float(np.max(np.abs(np.asarray(out[0]) - np.asarray(ref[0]))))
                ok = (detail['best_index'][0] == detail['best_index'][1] and detail['best_cost_rel'] < 1e-5
                      and np.allclose(out[3], ref[3], rtol=1e-4, atol=1e-4) and np.allclose(out[0], ref[0], rtol=1e-4, atol=1e-3))
                if method == 'cem_mppi':
                    ok = ok and np.allclose(out[7], ref[7], rtol=1e-4, atol=1e-5)
                    detail['sigma_maxabs'] = float(np.max(np.abs(np.asarray(out[7]) - np.asarray(ref[7]))))
                del solo
            mine = torch.tensor(np.concatenate([np.asarray(out[3], np.float32), np.asarray(out[0], np.float32)]), device='cuda')
            allr = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(allr, mine)
            identical = all(torch.equal(allr[0], a) for a in allr)
            # ---- latency of the boundary call ----
            best = np.asarray(out[3], np.float32)
            lat = np.zeros(args.ticks)
            for i in range(20 + args.ticks):
                if i == 20:
                    dist.barrier(); torch.cuda.synchronize()
                ctrl.with_newkey()
                t0 = time.perf_counter()
                o = tick(ctrl, sc, best, sigma)
                t1 = time.perf_counter()
                best = np.asarray(o[3], np.float32)
                if i >= 20:
                    lat[i - 20] = t1 - t0
            tot = torch.tensor([lat.sum()], dtype=torch.float64, device='cuda')
            dist.all_reduce(tot, op=dist.ReduceOp.MAX)
            # ---- what the exchange itself costs: in-kernel wait of every rank (its own GPU clock), 60 ticks ----
            waits = None
            if exchange == 'peer':
                ctrl.debug_peer_wait(True)
                w = np.zeros(60)
                for i in range(60):
                    ctrl.with_newkey()
                    o = tick(ctrl, sc, best, sigma)
                    best = np.asarray(o[3], np.float32)
                    a, b = ctrl.debug_peer_wait(True)
                    w[i] = (b - a) * 1e-3
                ctrl.debug_peer_wait(False)
                wt = torch.tensor(w, dtype=torch.float64, device='cuda')
                allw = [torch.zeros_like(wt) for _ in range(world)]
                dist.all_gather(allw, wt)
                allw = torch.stack(allw).cpu().numpy()                       # [rank][tick] us
                # the rank that arrives LAST waits only for the transfer; the one that arrives first also waits for the skew
                waits = {'exchange_us_p50': float(np.median(allw.min(axis=0))), 'skew_plus_exchange_us_p50': float(np.median(allw.max(axis=0))),
                         'skew_plus_exchange_us_p99': float(np.percentile(allw.max(axis=0), 99))}
            if rank == 0:
                line = {'what': 'sharded_controller', 'exchange': exchange, 'n_gpus': world, 'robot': robot, 'method': method,
                        'param': param, 'adaptive': adaptive, 'N': N, 'parity_ok': bool(ok), 'ranks_bit_identical': bool(identical),
                        'detail': detail, 'ticks': args.ticks, 'p50_us': float(np.percentile(lat, 50) * 1e6),
                        'p99_us': float(np.percentile(lat, 99) * 1e6),
                        'rollout_steps_per_s': args.ticks * N * ctrl.horizon / float(tot.item()),
                        'elapsed_s': round(time.perf_counter() - t_start, 1)}
                if waits:
                    line['in_kernel_exchange'] = waits
                print(json.dumps(line), flush=True)
                lines.append(line)
            del ctrl                     # (NCCL-exchange handles release their communicator with ncclCommAbort)
            dist.barrier()
    if rank == 0 and args.out:
        with open(args.out, 'w') as f:
            for l in lines:
                f.write(json.dumps(l) + '\n')
        bad = [l for l in lines if not (l['parity_ok'] and l['ranks_bit_identical'])]
        print(f'{len(lines)} cases, {len(bad)} failed', flush=True)
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    # torch's own NCCL teardown has been seen to hang after library-owned communicators (exchange = nccl) were destroyed
    # in the same process; all results are out by now, so leave without it.
    os._exit(0)


if __name__ == '__main__':
    main()
