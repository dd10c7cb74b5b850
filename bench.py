#!/usr/bin/env python
"""bench.py - telemetry samples/s through join + windowed statistics (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One "step" = one pass of the hot path over one batch of synthetic samples:
(node,pid)->pod join + per-pod / per-GPU / per-node 1-min tumbling mean / max / p95.
Workload at N=1: BASELINE.json configs[1] (1k nodes x 8 GPUs x 1 h @ 1 s = 28.8 M samples,
10k pods).  At N>1 every rank holds its own configs[1]-sized shard of nodes (weak scaling, the
node-sharding of configs[3]) and the step ends with the NCCL all_gather of the per-shard
result tables.

  value  device-resident throughput: the batch is already in HBM, CUDA events around the K steps
         (the K-step loop is repeated inside the two events until the timed region is >= 100 ms of
         device time, so that the clock sampler sees it; ms_per_step = elapsed / (K * repeats));
  e2e    the same job through the public API from pinned HOST columns: chunked
         cudaMemcpyAsync overlapped with the kernels, table build, D2H of every result table;
  roofline  every kernel of a step timed per launch with CUDA events in a separate pass; the block
         names the kernel that takes longest, gives each kernel's own algorithmic bytes and the whole
         step at 16 algorithmic bytes per sample, all against MEASURED_PEAKS.json hbm_gbs;
  cpu_baseline  the pure-Python oracle (the reference's algorithmic style; the Python-2
         reference itself cannot run) on a bounded slice of the same trace, 1 core.

``--impl reference`` times only that CPU arm (rank 0; other ranks exit).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = 'telemetry samples/sec (join + windowed stats)'
UNIT = 'samples/s'
QS = (0.95,)
KINDS = ('pod', 'gpu', 'node')
STEP_SEC, N_WINDOWS = 60, 60
ALGO_BYTES_PER_SAMPLE = 16


def workload_name(n_nodes):
    return ('configs[1]: %d nodes x 8 GPUs x 3600 s @ 1 Hz, %d pods, per-pod+per-GPU+per-node 1-min tumbling '
            'mean/max/p95 of sm_util' % (n_nodes, n_nodes * 10))


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            return float(json.load(f)['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md 6.65 TB/s)'


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.gpu), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            f = [x.strip() for x in r.split(',')]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'samples': len(sm), 'reasons': sorted(reasons)}


# ------------------------------------------------------------------------------------------
# CPU arm: the oracle's pure-Python statement of the path on a bounded slice
# ------------------------------------------------------------------------------------------

def cpu_arm_once(tr, n_nodes_slice):
    """Join + the three tables with plain Python loops on the first n_nodes_slice nodes."""
    from oracle import window_stats as ws
    n = n_nodes_slice * tr.gpus_per_node * tr.seconds
    t = time.perf_counter()
    table = ws.build_table(tr.keys, tr.pod_ids)
    for kind, nk in ((ws.KEY_POD, tr.n_pods), (ws.KEY_GPU, tr.n_gpus), (ws.KEY_NODE, tr.n_nodes)):
        ws.window_stats_py(tr.ts[:n], tr.gpu_gid[:n], tr.pid[:n], tr.util[:n], table, key_kind=kind, metric=0,
                           gpus_per_node=tr.gpus_per_node, t0=tr.t0, step=STEP_SEC, width=STEP_SEC,
                           n_windows=N_WINDOWS, n_keys=nk, qs=QS)
    return n, time.perf_counter() - t


def cpu_rate_probe(tr):
    n, dt = cpu_arm_once(tr, 1)
    return n / dt


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return 0
    from kube_gpu_stats_b200 import synth
    n_nodes = 16
    tr = synth.cluster_trace(n_nodes, 8, 1, 3600, n_tenancies=n_nodes * 10, seed=0, name='cfg2-slice')
    rate = cpu_rate_probe(tr)
    budget = 150.0 / max(1, args.steps + args.warmup)         # whole run ~2.5 min
    nodes = int(max(1, min(n_nodes, rate * budget / (8 * 3600))))
    for _ in range(args.warmup):
        cpu_arm_once(tr, nodes)
    t_all, n_all = 0.0, 0
    for _ in range(args.steps):
        n, dt = cpu_arm_once(tr, nodes)
        t_all += dt
        n_all += n
    value = n_all / t_all
    sample = '%d nodes x 8 GPUs x 3600 s (%d samples) of the configs[1] trace per step' % (nodes, nodes * 8 * 3600)
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * t_all / args.steps,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'u8 in / u64 + f64 out',
        'data': 'synthetic', 'config': {'workload': workload_name(1000), 'sample': sample},
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': 1, 'kind': 'port', 'sample': sample,
                         'host_cores': os.cpu_count(),
                         'note': 'pure-Python oracle restating the reference style (dict probes, list append, '
                                 'sum/len); the Python-2 reference cannot run on this image; it is single-threaded'},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------

def run_b200(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        os.environ.setdefault('MASTER_PORT', '29511')
    if not torch.cuda.is_available():
        raise SystemExit('bench.py --impl b200 needs a CUDA device (there is no CPU fallback)')
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))

    from kube_gpu_stats_b200 import build, synth
    from kube_gpu_stats_b200.engine import Engine
    from kube_gpu_stats_b200 import sharding
    build.build_lib()

    n_nodes = args.nodes
    tr = synth.cluster_trace(n_nodes, 8, 1, 3600, n_tenancies=n_nodes * 10, seed=rank, name='cfg2')
    n = tr.n
    eng = Engine(local)
    if world > 1:
        eng.set_packed_exchange('compact')    # compact tables on the wire (config.exchange_bytes_per_cell)
    # pinned host columns (the e2e leg copies from these)
    host = [eng.pinned(n) for _ in range(4)]
    for dst, src in zip(host, (tr.ts, tr.gpu_gid, tr.pid, tr.util)):
        dst[:] = src
    spec = eng.make_spec(kinds=KINDS, metric='sm', gpus_per_node=8, t0=tr.t0, step=STEP_SEC, n_windows=N_WINDOWS,
                         n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=QS, total_samples_hint=n)
    eng.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    eng.ingest(*host)
    eng.sync()

    stream = torch.cuda.ExternalStream(eng.stream('compute'), device=torch.device('cuda', local))
    gather, exchange = None, None
    if world > 1:
        # The tables travel through the library's own exchange (kgs_exchange_push: copy-engine pushes into
        # peer-mapped memory + one barrier kernel, no SMs taken from the next job's kernels); torch supplies the
        # peer-mapped allocation.  kgs_exchange_nccl (one ncclAllGather issued by the library on its own
        # communicator) is the transport when symmetric memory cannot be set up, or with KGS_EXCHANGE=nccl.
        with torch.cuda.stream(stream):
            eng.window_stats_device(spec)
        eng.sync()
        slot_bytes = eng.exchange_region()[1]
        try:
            if os.environ.get('KGS_EXCHANGE', 'peer') != 'peer':
                raise RuntimeError('KGS_EXCHANGE asks for NCCL')
            gather = sharding.PeerExchange(eng, world, rank, slot_bytes, self_copy=False,   # own tables stay in the result region
                                           multicast=os.environ.get('KGS_EXCHANGE_MULTICAST', '0') == '1')
            exchange = ('kgs_exchange_push: copy-engine pushes of the compact tables into every peer over NVLink '
                        '(peer-mapped symmetric memory) + one barrier kernel, overlapped with the next step')
            if gather.multicast:
                exchange = ('kgs_exchange_push: ONE copy-engine push of the compact tables to the NVSwitch multicast address '
                            '(replicated into every rank\'s buffer) + one barrier kernel, overlapped with the next step')
        except Exception as exc:                 # noqa: BLE001
            sys.stderr.write('rank %d: peer exchange not available (%s); using the NCCL all_gather\n' % (rank, exc))
            ok = torch.tensor([0], device='cuda')
        else:
            ok = torch.tensor([1], device='cuda')
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok) == 0:
            ids = [eng.nccl_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(ids, src=0)
            eng.nccl_init(ids[0], world, rank)
            gather = sharding.NcclExchange(eng, world, slot_bytes)
            exchange = 'kgs_exchange_nccl: one ncclAllGather of the packed tables (library-owned communicator), overlapped with the next step'

    def step():
        # N > 1: the exchange of step i's tables runs on the library's exchange stream while step i+1
        # computes (double-buffered result regions, ordering inside the library); everything is joined
        # before the closing barrier
        eng.window_stats_device(spec)
        if gather is not None:
            gather.push(want_view=False)

    def barrier():
        if gather is not None:
            eng.exchange_sync()
        if world > 1:
            dist.barrier()
        eng.sync()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        for _ in range(max(3, args.warmup)):
            step()
        barrier()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        # how often the K-step loop has to run for >= 100 ms of device time (same count on every rank)
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record(stream)
        for _ in range(args.steps):
            step()
        if gather is not None:
            eng.exchange_join()
        p1.record(stream)
        barrier()
        probe_ms = torch.tensor([p0.elapsed_time(p1)], dtype=torch.float64, device='cuda')
        if world > 1:
            dist.all_reduce(probe_ms, op=dist.ReduceOp.MIN)
        repeats = int(max(1, min(10000, np.ceil(args.min_timed_ms / max(float(probe_ms), 1e-3)))))
        launches0 = eng.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(repeats):
            for _ in range(args.steps):
                step()
        if gather is not None:
            eng.exchange_join()          # the last exchanges belong to the timed region
        e1.record(stream)
        barrier()
        ms_total = e0.elapsed_time(e1) / repeats       # per K steps
        launches = (eng.launch_count() - launches0) // repeats

        # ---- e2e: host columns -> result tables on the host, every step -----------------
        n_chunks = 8
        cuts = [(n * i // n_chunks) // 128 * 128 for i in range(n_chunks)] + [n]
        h2d_bytes = n * 16 + len(tr.keys) * 12

        def e2e_step():
            eng.build_join(tr.keys, tr.pod_ids, tr.n_pods)
            eng.stats_begin(spec)
            for a, b in zip(cuts[:-1], cuts[1:]):
                eng.ingest(host[0][a:b], host[1][a:b], host[2][a:b], host[3][a:b])
                eng.stats_accumulate()
            eng.stats_finalize()
            if gather is not None:
                gather.push()
                eng.exchange_join()
            return eng.fetch_all(KINDS)          # one D2H of the whole result region into pinned memory

        for _ in range(2):
            e2e_step()
        barrier()
        e2e_steps = max(3, min(args.steps, 10))
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            res = e2e_step()
        barrier()
        e2e_s = time.perf_counter() - t0
        d2h_bytes = eng.result_region()[2]
        xbytes_per_rank = eng.exchange_region()[1] if world > 1 else 0
        if rank == 0:
            clocks = sampler.stop()

        # ---- copy-only leg: the same pinned chunks through cudaMemcpyAsync, no kernels ------------
        # (what the host side of this box can deliver to N GPUs at once; explains e2e at N > 1)
        def copy_only():
            for a, b in zip(cuts[:-1], cuts[1:]):
                eng.ingest(host[0][a:b], host[1][a:b], host[2][a:b], host[3][a:b])
            eng.sync()
        copy_only()
        barrier()
        tc = time.perf_counter()
        for _ in range(3):
            copy_only()
        barrier()
        copy_s = (time.perf_counter() - tc) / 3

        # ---- per-kernel pass (events around every launch; not the throughput pass) -------
        eng.ingest(*host)
        eng.sync()
        eng.set_profiling(True)
        eng.reset_kernel_times()
        prof_steps = 10
        for _ in range(prof_steps):
            eng.window_stats_device(spec)
        eng.sync()
        ktimes = eng.kernel_times()
        eng.set_profiling(False)

    t = torch.tensor([ms_total, e2e_s, copy_s], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_s, copy_s = float(t[0]), float(t[1]), float(t[2])
    total_samples = n * world

    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        # every kernel of a step: measured time, its own algorithmic bytes, and the DRAM
        # traffic of one launch from the committed ncu capture (profiles/*_traffic.json) where there is one
        n_q = len(QS)
        cells = {'pod': tr.n_pods * N_WINDOWS, 'gpu': tr.n_gpus * N_WINDOWS, 'node': tr.n_nodes * N_WINDOWS}
        n_cells = sum(cells.values())
        cell_io = 20 + 28 + 8 * n_q                       # accumulators in, count/sum/max/mean/pct out
        algo = {
            'accum': ALGO_BYTES_PER_SAMPLE * n,                                            # the four sample columns, once
            'finalize_groups': n + cells['gpu'] * 16 + n_cells * cell_io,                  # value log once, one run descriptor per gpu cell
            'finalize_cells': n_cells * 1,                                                 # the done flags (cells it has to finish itself: data dependent)
            'finalize_wide': 16,
        }
        traffic_file = {'accum': 'accum_traffic.json', 'finalize_groups': 'finalize_groups_traffic.json',
                        'finalize_cells': 'finalize_traffic.json'}

        def traffic_of(name):
            path = os.path.join(ROOT, 'profiles', traffic_file.get(name, '-'))
            try:
                return json.load(open(path)).get('dram_bytes_per_launch')
            except Exception:
                return None
        kernels = {}
        for name, (ms, cnt) in ktimes.items():
            if not cnt:
                continue
            avg = ms / cnt
            ent = {'avg_launch_ms': avg, 'launches_per_step': cnt / prof_steps, 'ms_per_step': ms / prof_steps,
                   'algorithmic_bytes_per_launch': algo.get(name), 'traffic': traffic_of(name)}
            if algo.get(name):
                ent['achieved'] = algo[name] / (avg * 1e-3) / 1e9
                ent['frac'] = ent['achieved'] / peak
            kernels[name] = ent
        step_kernel_ms = sum(e['ms_per_step'] for e in kernels.values())
        top = max(kernels, key=lambda k: kernels[k]['ms_per_step'])
        step_ms = ms_total / args.steps
        step_achieved = ALGO_BYTES_PER_SAMPLE * n / (step_ms * 1e-3) / 1e9
        # bounded CPU baseline on the same trace
        cpu = None
        crossover = None
        if not args.no_cpu_baseline and world == 1:
            # trace size from which the GPU path, transfers included, beats the CPU path (north_star)
            sys.path.insert(0, os.path.join(ROOT, 'tools'))
            import crossover as xo
            xrows, xn = xo.measure(sizes=(8, 32, 128, 512, 2048, 8192))
            crossover = {'gpu_faster_from_samples': xn,
                         'points': [{'samples': n_, 'gpu_e2e_ms': g_ * 1e3, 'cpu_ms': c_ * 1e3 if c_ else None} for n_, g_, c_ in xrows],
                         'note': 'GPU = table build + pinned H2D + join + 3 tables with p95 + D2H; CPU = pure-Python oracle, 1 core'}
        decode = None
        if not args.no_cpu_baseline and world == 1:
            # the reference's own input: text of a Prometheus answer, decoded on the GPU (row a1)
            import decode_bench
            decode = decode_bench.measure(series=8 * n_nodes, points=3600, cpu_series=256, reps=3)
        if not args.no_cpu_baseline:
            rate = cpu_rate_probe(tr)
            nodes = int(max(1, min(n_nodes, rate * 15.0 / (8 * 3600))))
            cn, cdt = cpu_arm_once(tr, nodes)
            cpu = {'value': cn / cdt, 'unit': UNIT, 'cores': 1, 'kind': 'port', 'host_cores': os.cpu_count(),
                   'sample': 'first %d nodes (%d samples) of the same trace, all three tables + p95' % (nodes, cn)}
        line = {
            'metric': METRIC, 'value': total_samples * args.steps / (ms_total * 1e-3), 'unit': UNIT,
            'n_gpus': world, 'steps': args.steps, 'warmup': max(3, args.warmup),
            'ms_per_step': ms_total / args.steps, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'u8 in / u64 + f64 out', 'data': 'synthetic',
            'config': {'workload': workload_name(n_nodes), 'samples_per_gpu': n, 'bytes_per_sample': 16,
                       'l2': 'input per step is %.0f MB, larger than the 126 MB L2 (no flush needed)' % (n * 16 / 1e6),
                       'parallelism': 'node-sharded x%d' % world, 'collective': exchange,
                       'exchange_bytes_per_rank_per_step': xbytes_per_rank,
                       'exchange_bytes_per_cell': (xbytes_per_rank / n_cells) if world > 1 else None,
                       'timed_repeats': repeats, 'timed_region_ms': ms_total * repeats},
            'clocks': clocks,
            'e2e': {'value': total_samples * e2e_steps / e2e_s, 'unit': UNIT, 'h2d_bytes_per_step': h2d_bytes,
                    'd2h_bytes_per_step': d2h_bytes, 'steps': e2e_steps, 'ms_per_step': 1e3 * e2e_s / e2e_steps,
                    'timing': 'host wall clock around blocking API calls, sync both sides',
                    'copy_only': {'ms_per_step': 1e3 * copy_s, 'aggregate_h2d_gbs': world * n * 16 / copy_s / 1e9,
                                  'note': 'the same 8 pinned chunks per rank through cudaMemcpyAsync with no kernels: '
                                          'what the host side delivers to %d GPU(s) at once' % world}},
            'gpu_launches': int(launches),
            'roofline': {'bound': 'hbm', 'kernel': 'k_' + top, 'achieved': kernels[top].get('achieved'), 'peak': peak,
                         'unit': 'GB/s', 'frac': kernels[top].get('frac'), 'traffic': kernels[top]['traffic'],
                         'peak_source': peak_src, 'avg_launch_ms': kernels[top]['avg_launch_ms'],
                         'algorithmic_bytes_per_launch': kernels[top]['algorithmic_bytes_per_launch'],
                         'frac_of_nominal_8TBs': (kernels[top].get('achieved') or 0) / 8000.0,
                         'step': {'algorithmic_bytes': ALGO_BYTES_PER_SAMPLE * n, 'ms': step_ms, 'achieved': step_achieved,
                                  'frac': step_achieved / peak, 'frac_of_nominal_8TBs': step_achieved / 8000.0,
                                  'note': '16 B per sample over the driver-timed step (all kernels, launch gaps%s)'
                                          % (', table exchange' if world > 1 else '')},
                         'kernels': kernels,
                         'kernel_ms_sum_per_step': step_kernel_ms,
                         'note': 'per-kernel times from a separate pass with CUDA events around every launch on the '
                                 'launching stream; kernel = the one with the largest share of a step'},
            'cpu_baseline': cpu,
            'crossover': crossover,
            'decode': decode,
        }
        # keep the result honest: the last e2e step's tables must not be empty
        assert int(res['node']['count'].sum()) == n, 'e2e tables do not cover the batch'
        print(json.dumps(line))
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--nodes', type=int, default=1000, help='nodes per GPU shard (configs[1] = 1000)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--min-timed-ms', type=float, default=100.0, help='repeat the K-step loop until the timed region is this long')
    args = ap.parse_args()
    if args.impl == 'reference':
        return run_reference(args)
    world = int(os.environ.get('WORLD_SIZE', '1'))
    if args.gpus > 1 and world == 1:
        # convenience: relaunch under torchrun when called plainly with --gpus N
        cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(args.gpus),
               '--master-addr', '127.0.0.1', '--master-port', '29533', os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    return run_b200(args)


if __name__ == '__main__':
    sys.exit(main())
