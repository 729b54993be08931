#!/usr/bin/env python
# -*- coding: utf-8 -*-
"""Flagship benchmark: fwd+bwd of DistributedDotProductAttn (BASELINE.json config 4 per-rank shape).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--T total_len]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Model: 16 heads, d_model = 1024, batch 1, bf16, T/N = 8192 rows per GPU (T = 8192 * N; N = 8 is the named
T = 65536 config).  Synthetic random inputs, random-init weights.  One step = forward + backward of the whole
layer (4 projections, K Q^T, scale+mask+softmax, attn V and all their gradients).

Prints ONE JSON line (rank 0).  ``value`` = ms per fwd+bwd step, device-timed with CUDA events around K steps,
max over ranks.  ``e2e`` = the same through the public API including, every step, the pinned-host -> device copy
of the step's activations and target and the device -> host read of the loss.

``--impl reference`` runs the unmodified reference from baseline/_ref (Horovod/mpi4py replaced by the
torch.distributed shim in baseline/shim) on the same shapes in its native dtype (fp32).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))

HEADS = 16
D_MODEL = 1024
ROWS_PER_GPU = 8192


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', type=str, default='ours', choices=['ours', 'reference'])
    ap.add_argument('--T', type=int, default=0, help='total sequence length (default 8192 * gpus)')
    ap.add_argument('--heads', type=int, default=HEADS)
    ap.add_argument('--d-model', type=int, default=D_MODEL)
    ap.add_argument('--offset', type=int, default=1024)
    ap.add_argument('--dtype', type=str, default='bf16', help='ours: bf16 | fp16 | fp32 (fp32 storage, TF32 math)')
    ap.add_argument('--graph', type=str, default='auto', choices=['auto', 'off'],
                    help='ours, 1 GPU: replay the whole step from a CUDA graph (utils.graphs.GraphedStep)')
    ap.add_argument('--compute-dtype', type=str, default='', choices=['', 'bf16', 'fp16'],
                    help='ours, with --dtype fp32: mixed precision - fp32 parameters / activations, 16-bit math on the '
                         'slab-free kernels (DistributedDotProductAttn(compute_dtype=...))')
    ap.add_argument('--ref-autocast', type=str, default='none', choices=['none', 'bf16'],
                    help='reference arm: run the stock module inside torch.autocast(cuda, bfloat16)')
    ap.add_argument('--mask', type=str, default='zeros', choices=['none', 'zeros', 'causal', 'random'],
                    help="attention mask: the reference example's all-False mask (default), none, causal, or 30%% random")
    ap.add_argument('--mask-refresh', type=int, default=0,
                    help='build a NEW mask tensor every this many steps (0 = one mask for the whole run), so that '
                         'the packed-mask cache of the slab-free path misses and the byte mask is read again')
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of this process's GPU during the timed region (NVML)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.stop_flag = threading.Event()
        self.sm, self.reasons, self.max_mhz, self.power = [], set(), None, []
        self.ok = False
        self.t_begin = self.t_end = None
        self.wake = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, 'nvmlClocksThrottleReasonHwSlowdown', 0x8): 'hw_slowdown',
            getattr(nv, 'nvmlClocksThrottleReasonHwThermalSlowdown', 0x40): 'hw_thermal_slowdown',
            getattr(nv, 'nvmlClocksThrottleReasonSwThermalSlowdown', 0x20): 'sw_thermal_slowdown',
            getattr(nv, 'nvmlClocksThrottleReasonSwPowerCap', 0x4): 'sw_power_cap',
            getattr(nv, 'nvmlClocksThrottleReasonHwPowerBrakeSlowdown', 0x80): 'hw_power_brake',
        }
        # The thread is started BEFORE the warm-up steps (thread start-up and NVML's first calls stay out of the
        # timed region); only samples taken between mark_begin() and mark_end() are reported.
        while not self.stop_flag.is_set():
            try:
                live = self.t_begin is not None and self.t_end is None
                sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                if live and self.t_end is None:
                    self.sm.append(sm)
                    for bit, name in names.items():
                        if r & bit:
                            self.reasons.add(name)
                    self.power.append(pw)
            except Exception:
                pass
            self.wake.wait(0.02 if (self.t_begin is not None and self.t_end is None) else 0.05)
            self.wake.clear()

    def mark_begin(self):
        self.t_begin = time.time()
        self.wake.set()   # sample now, not after the idle-period sleep

    def mark_end(self):
        self.t_end = time.time()

    def summary(self):
        if not self.ok or not self.sm:
            return {'sm_mhz': None, 'sm_max_mhz': self.max_mhz, 'reasons': ['unavailable']}
        s = sorted(self.sm)
        return {'sm_mhz': s[len(s) // 2], 'sm_max_mhz': self.max_mhz, 'reasons': sorted(self.reasons),
                'power_w_max': max(self.power) if self.power else None, 'samples': len(s)}


def physical_gpu_index(local_rank):
    vis = os.environ.get('CUDA_VISIBLE_DEVICES')
    if vis:
        try:
            return int(vis.split(',')[local_rank])
        except Exception:
            return local_rank
    return local_rank


def main():
    args = parse()
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if args.gpus != world and world == 1 and args.gpus > 1:
        # convenience: re-exec under torchrun when asked for N > 1 without a launcher
        os.execvp(sys.executable, [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1',
                                   '--nproc-per-node', str(args.gpus), '--master-addr', '127.0.0.1',
                                   '--master-port', '29577', os.path.abspath(__file__)] + sys.argv[1:])
    reference = args.impl == 'reference'
    if reference:
        ref_dir = os.path.join(ROOT, 'baseline', '_ref')
        if not os.path.exists(os.path.join(ref_dir, 'distributed_dot_product', 'module.py')):
            if rank == 0:
                print(json.dumps({'impl': 'reference',
                                  'unavailable': 'baseline/_ref is missing: run baseline/install_reference.sh'}))
            return 0
        # the reference package has the same import name as ours: make it (and the hvd/mpi4py shim) win
        sys.path[:] = [p for p in sys.path if os.path.abspath(p or '.') != ROOT]
        sys.path.insert(0, ref_dir)
        sys.path.insert(0, os.path.join(ROOT, 'baseline', 'shim'))
    import torch
    if not torch.cuda.is_available():
        if rank == 0:
            print(json.dumps({'impl': args.impl, 'unavailable': 'no CUDA device'}))
        return 0
    torch.cuda.set_device(local_rank % torch.cuda.device_count())
    dev = torch.device('cuda', torch.cuda.current_device())
    import torch.distributed as dist

    if reference:
        import horovod.torch as hvd
        hvd.init()
        torch.backends.cuda.matmul.allow_tf32 = True
        torch.backends.cudnn.allow_tf32 = True
        from distributed_dot_product.module import DistributedDotProductAttn
        import distributed_dot_product as pkg
        assert os.path.abspath(pkg.__file__).startswith(os.path.join(ROOT, 'baseline', '_ref')), pkg.__file__
        dtype = torch.float32
        count_launches = None
    else:
        from distributed_dot_product.utils import comm
        comm.init()
        from distributed_dot_product.module import DistributedDotProductAttn
        from distributed_dot_product import _native
        _native.load()
        dtype = {'bf16': torch.bfloat16, 'fp16': torch.float16, 'fp32': torch.float32}[args.dtype]
        count_launches = _native
    world = dist.get_world_size() if dist.is_initialized() else 1

    total_t = args.T if args.T else ROWS_PER_GPU * world
    tn = total_t // world
    torch.manual_seed(111)
    mixed = None
    if not reference and args.compute_dtype:
        if dtype != torch.float32:
            raise SystemExit('--compute-dtype goes with --dtype fp32')
        mixed = {'bf16': torch.bfloat16, 'fp16': torch.float16}[args.compute_dtype]
    model = DistributedDotProductAttn(args.d_model, num_heads=args.heads, offset=args.offset,
                                      **({'compute_dtype': mixed} if mixed is not None else {})).to(dev, dtype)
    if world > 1:
        for p in model.parameters():
            dist.broadcast(p.data, src=0)
    torch.manual_seed(222 + rank)
    # synthetic data: pinned host buffers (e2e path) and resident device copies (device-timed path)
    x_host = torch.rand(1, tn, args.d_model).to(dtype).pin_memory()
    y_host = torch.rand(1, tn, args.d_model).to(dtype).pin_memory()
    x_dev = x_host.to(dev)
    y_dev = y_host.to(dev)
    def make_mask():
        if args.mask == 'none':
            return None
        if args.mask == 'zeros':
            return torch.zeros(1, tn, total_t, device=dev).bool()  # the reference example's all-False mask
        if args.mask == 'causal':
            rows = torch.arange(rank * tn, (rank + 1) * tn, device=dev).view(1, tn, 1)
            return torch.arange(total_t, device=dev).view(1, 1, total_t) > rows
        g = torch.Generator(device=dev)
        g.manual_seed(333 + rank)
        return torch.rand(1, tn, total_t, device=dev, generator=g) < 0.3

    if reference and args.mask == 'none':
        raise SystemExit('the reference module needs a bool mask (module.py:66)')
    state = {'mask': make_mask(), 'n': 0}
    loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
    autocast = reference and args.ref_autocast == 'bf16'

    def step(x, y):
        if args.mask_refresh and state['n'] and state['n'] % args.mask_refresh == 0:
            state['mask'] = make_mask()   # a new tensor object: nothing cached for the old one can be reused
        state['n'] += 1
        mask = state['mask']
        for p in model.parameters():
            p.grad = None
        if autocast:
            with torch.autocast('cuda', dtype=torch.bfloat16):
                out = model(x, x, x, mask)
        else:
            out = model(x, x, x, mask)
        loss = torch.nn.functional.mse_loss(out.float(), y.float())
        loss.backward()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # L2 flush between timed steps: the slab-free attention path has a working set (projections, gradients)
    # well below the 126 MB L2, so every timed step starts from a cache that a 256 MB write has just evicted.
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    step_log = []   # per-step times of this rank, one list per timed region (device-timed, end-to-end)

    def timed_steps(fn):
        """K steps, each bracketed by its own CUDA event pair, L2 flushed in between (outside the pairs)."""
        pairs = []
        for _ in range(args.steps):
            flush_buf.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            pairs.append((a, b))
        barrier()
        per_step = [a.elapsed_time(b) for a, b in pairs]
        step_log.append([round(v, 3) for v in per_step])
        return sum(per_step) / args.steps

    result = {'impl': args.impl}
    attention_path = None
    self_check = None
    try:
        if not reference:
            # Self-check of the slab-free attention kernels against the slab path on this very configuration
            # (untimed): if the losses disagree or a peer wait timed out, the bench falls back to the slab path.
            import distributed_dot_product.ops.attention as att
            from distributed_dot_product.parallel.runtime import check_status
            attention_path = 'slab'
            # the slab path keeps P and dS ([heads, T/N, T] each) plus transients: only check against it when it fits
            slab_bytes = 3.5 * args.heads * tn * total_t * 2
            fits = slab_bytes < 0.8 * torch.cuda.mem_get_info()[0]
            if att.FLASH_TRAIN and dtype in (torch.bfloat16, torch.float16) and not fits:
                attention_path = 'slab-free (flash)'
                self_check = {'skipped': 'the slab path needs ~%.0f GB at this shape' % (slab_bytes / 1e9)}
            elif att.FLASH_TRAIN and dtype in (torch.bfloat16, torch.float16):
                # element-wise: output and input gradient of the slab-free path against the slab path
                def probe():
                    xg = x_dev.clone().requires_grad_(True)
                    for p in model.parameters():
                        p.grad = None
                    out = model(xg, xg, xg, state['mask'])
                    torch.nn.functional.mse_loss(out.float(), y_dev.float()).backward()
                    return out.detach().float(), xg.grad.float()

                ok = 0.0
                try:
                    o_f, g_f = probe()
                    check_status()
                    att.FLASH_TRAIN = False
                    o_s, g_s = probe()
                    e_out = ((o_f - o_s).abs().max() / o_s.abs().max().clamp_min(1e-12)).item()
                    e_grad = ((g_f - g_s).abs().max() / g_s.abs().max().clamp_min(1e-30)).item()
                    self_check = {'max_rel_err_out': e_out, 'max_rel_err_x_grad': e_grad, 'against': 'slab path'}
                    ok = 1.0 if (e_out <= 3e-2 and e_grad <= 3e-2) else 0.0
                    del o_f, g_f, o_s, g_s
                except RuntimeError as e:  # watchdog / peer timeout: keep the proven path
                    sys.stderr.write('slab-free attention self-check failed: %s\n' % e)
                    self_check = {'error': str(e).splitlines()[0][:200]}
                flag = torch.tensor([ok], device=dev)
                if world > 1:  # every rank gets here, also the one that raised
                    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                att.FLASH_TRAIN = bool(flag.item() > 0)
                attention_path = 'slab-free (flash)' if att.FLASH_TRAIN else 'slab (self-check rejected slab-free path)'
                torch.cuda.empty_cache()
            elif mixed is not None:
                attention_path = 'slab-free (flash), fp32 parameters and activations cast per call'
            elif dtype == torch.float32:
                attention_path = 'slab (fp32 storage, TF32 tensor-core math)'
        torch.cuda.reset_peak_memory_stats()  # the self-check's slab step must not count as this path's peak
        sampler = ClockSampler(physical_gpu_index(local_rank))
        sampler.start()
        for _ in range(max(args.warmup, 3)):
            step(x_dev, y_dev)
        barrier()
        # ---------------- device-timed region
        # one GPU: the step is replayed from a CUDA graph (one launch instead of ~26); the launch count is taken
        # from an eager step
        graphed = None
        use_graph = (not reference and world == 1 and args.graph == 'auto' and not args.mask_refresh)
        run_dev = lambda: step(x_dev, y_dev)  # noqa: E731
        per_step_launches = None
        if use_graph:
            try:
                from distributed_dot_product.utils.graphs import GraphedStep
                l0 = count_launches.launch_count
                step(x_dev, y_dev)
                per_step_launches = count_launches.launch_count - l0
                graphed = GraphedStep(lambda: step(x_dev, y_dev))
                eager_loss = step(x_dev, y_dev).item()
                if abs(graphed().item() - eager_loss) > 1e-3 * abs(eager_loss) + 1e-12:
                    raise RuntimeError('graph replay and eager step disagree')
                run_dev = graphed
            except Exception as e:  # noqa: BLE001 - any capture problem: time the eager step
                sys.stderr.write('CUDA graph capture failed, timing the eager step: %s\n' % e)
                graphed = None
                torch.cuda.synchronize()
        l0 = count_launches.launch_count if count_launches else 0
        sampler.mark_begin()
        ms = timed_steps(run_dev)
        sampler.mark_end()
        launches = (count_launches.launch_count - l0) if count_launches else None
        if graphed is not None:
            launches = per_step_launches * args.steps
        sampler.stop_flag.set()
        sampler.join()
        # ---------------- end-to-end region: H2D of the step inputs + D2H of the loss, every step.
        # Input pipeline (both arms): two device buffers; the inputs of step i + 1 cross PCIe on a copy stream while
        # step i computes (every step still copies its own inputs from pinned memory inside the timed region; the
        # first timed step pays its copy in full, the last one prefetches a batch nobody uses).
        copy_stream = torch.cuda.Stream()
        bufs = [(torch.empty_like(x_dev), torch.empty_like(y_dev)) for _ in range(2)]
        ready = [None, None]   # copy of the buffer pair finished (copy stream)
        done = [None, None]    # last step that read the buffer pair finished (compute stream)
        pipe = {'i': 0}
        run_pair = [lambda: step(*bufs[0]), lambda: step(*bufs[1])]
        if graphed is not None:
            try:
                run_pair = [GraphedStep(run_pair[0]), GraphedStep(run_pair[1])]
            except Exception as e:  # noqa: BLE001
                sys.stderr.write('CUDA graph capture failed for the end-to-end step: %s\n' % e)
                torch.cuda.synchronize()

        def prefetch(i):
            if done[i & 1] is not None:
                copy_stream.wait_event(done[i & 1])   # step i - 2 has read this pair; step i - 1 is NOT waited for
            with torch.cuda.stream(copy_stream):
                bx, by = bufs[i & 1]
                bx.copy_(x_host, non_blocking=True)
                by.copy_(y_host, non_blocking=True)
                ready[i & 1] = torch.cuda.Event()
                ready[i & 1].record(copy_stream)

        def e2e_step():
            i = pipe['i']
            pipe['i'] += 1
            if ready[i & 1] is None:
                prefetch(i)          # pipeline fill
            torch.cuda.current_stream().wait_event(ready[i & 1])
            ready[i & 1] = None
            prefetch(i + 1)          # crosses PCIe while this step computes
            loss = run_pair[i & 1]()
            done[i & 1] = torch.cuda.Event()
            done[i & 1].record()
            loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
            torch.cuda.current_stream().synchronize()  # the caller reads the loss every step

        for _ in range(2):
            e2e_step()
        copy_stream.synchronize()
        ready[0] = ready[1] = None   # the timed region starts with an empty pipeline
        pipe['i'] = 0
        barrier()

        ms_e2e = timed_steps(e2e_step)
        t = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e = t.tolist()
        h2d = x_host.numel() * x_host.element_size() + y_host.numel() * y_host.element_size()
        flops_fwd = 4.0 * 2 * tn * args.d_model * args.d_model + 2 * 2.0 * tn * total_t * args.d_model
        result.update({
            'metric': ('DistributedDotProductAttn fwd+bwd ms/step (16 heads, d_model=1024, T/N=8192)' if not args.T
                       else 'DistributedDotProductAttn fwd+bwd ms/step (%d heads, d_model=%d, T=%d)'
                       % (args.heads, args.d_model, total_t)),
            'value': ms, 'unit': 'ms', 'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3),
            'ms_per_step': ms, 'higher_is_better': False, 'scaling': 'strong' if args.T else 'weak',
            'vs_baseline': None,
            'work_normalised_ms': ms / world,  # per-GPU attention work grows with N at fixed rows per GPU
            'dtype': (('bf16 autocast over fp32 weights (stock reference module under torch.autocast)' if autocast
                       else 'fp32 storage, tf32 matmul (reference native dtype)') if reference
                      else (('fp32 storage, %s math (mixed precision)' % args.compute_dtype) if args.compute_dtype
                            else 'fp32 storage, tf32 matmul' if args.dtype == 'fp32' else args.dtype)),
            'data': 'synthetic',
            'step_ms_rank0': {'device': step_log[0], 'e2e': step_log[1]},
            'config': {'model': 'DistributedDotProductAttn heads=%d d_model=%d' % (args.heads, args.d_model),
                       'global_batch': 1, 'seq_len': total_t, 'rows_per_gpu': tn,
                       'parallelism': 'sp%d (time axis split, all-gather-KV)' % world,
                       'offset': args.offset,
                       'e2e_pipeline': 'inputs of step i+1 prefetched on a copy stream while step i computes',
                       'cuda_graph': graphed is not None,
                       'mask': args.mask + (' (new tensor every %d steps)' % args.mask_refresh
                                            if args.mask_refresh else ' (one tensor for the run)'),
                       'l2': 'L2 flushed between timed steps (256 MB device write outside the per-step event pairs)',
                       'attention_path': attention_path if attention_path else 'reference (slab)',
                       'scaling_note': 'rows per GPU fixed; per-GPU attention FLOPs grow with N (T^2/N)'},
            'tflops_per_gpu': 3.0 * flops_fwd / (ms * 1e-3) / 1e12,
            'clocks': sampler.summary(),
            'e2e': {'value': ms_e2e, 'unit': 'ms', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': 4},
            'gpu_launches': launches if launches is not None else 0,
            'loss': float(loss_host.item()),
            'peak_mem_gb': torch.cuda.max_memory_allocated() / 1e9,
        })
        if self_check is not None:
            result['self_check'] = self_check
        if reference:
            result['gpu_launches_note'] = 'reference arm: library kernels only (cuBLAS/ATen/NCCL)'
    except torch.cuda.OutOfMemoryError as e:
        result = {'impl': args.impl, 'unavailable': 'CUDA out of memory at T=%d on %d GPU(s): %s'
                  % (total_t, world, str(e).split('.')[0])}
    except RuntimeError as e:
        if not (reference and autocast):
            raise
        # the stock reference cannot train under autocast: its hand-written backward runs outside the autocast
        # region and multiplies an fp32 gradient with a bf16 operand (functions.py:96)
        result = {'impl': args.impl, 'unavailable': 'stock reference under torch.autocast(bfloat16) raises in backward: %s'
                  % str(e).splitlines()[0][:160]}
    if rank == 0:
        print(json.dumps(result))
    if world > 1:
        try:
            dist.barrier()
            if not reference:
                from distributed_dot_product.utils import comm
                comm.shutdown()
            else:
                dist.destroy_process_group()
        except Exception:
            pass
    return 0


if __name__ == '__main__':
    sys.exit(main())
