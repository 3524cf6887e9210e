#!/usr/bin/env python
"""Headline benchmark: mel frames / second of ForwardTacotron.generate() at batch 64 x 200 phonemes
(BASELINE.json configs[1]) on N B200s, one process per GPU, sentence-sharded (weak scaling: every
rank synthesises its own 64 x 200 batch; no data-path collective).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Prints ONE JSON line (rank 0).  `value` = valid mel frames of all ranks / max-over-ranks device time
with token ids already resident in HBM; `e2e` = the same through host buffers (pinned H2D of the token
ids + D2H of mel_post inside the timed region).  `--impl reference` times the reference's CPU forward
(the oracle port of it: the Python reference itself cannot travel to the GPU box) on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = 'mel_frames_per_sec'
UNIT = 'frames/s'
SEED = 1234
# BASELINE.json configs.  'weak': every rank runs `batch` sentences (the driver's scaling run);
# 'strong': `batch` is the GLOBAL batch, sentence-sharded over the ranks (configs 4 and 5 are defined that way).
WORKLOADS = {
    'config2': dict(model='forward_tacotron', batch=64, T=200, alpha=1.0, scaling='weak', durations='fallback',
                    name='ForwardTacotron singlespeaker generate()'),
    'config3': dict(model='fast_pitch', batch=64, T=200, alpha=1.0, scaling='weak', durations='spread',
                    name='FastPitch (FFT-block variant) generate()'),
    'config4': dict(model='multi_forward_tacotron', batch=128, T=200, alpha=1.0, scaling='strong', durations='fallback',
                    name='MultiForwardTacotron generate() with 256-d speaker embedding'),
    'config5': dict(model='forward_tacotron', batch=32, T=4096, alpha=0.8, scaling='strong', durations='fallback',
                    name='ForwardTacotron long-article generate()'),
}


def workload_config(wl, durations, world):
    """The `config` object both arms print (same dict -> the driver's same_config check passes)."""
    w = WORKLOADS[wl]
    per = f"{w['batch']}x{w['T']} phonemes per GPU" if w['scaling'] == 'weak' else \
        f"{w['batch']}x{w['T']} phonemes sentence-sharded over the GPUs"
    return {'workload': f"{w['name']}, {per}, alpha={w['alpha']}, random-init, durations={durations}",
            'baseline_config': wl, 'model': w['model'], 'batch': w['batch'], 'tokens': w['T'], 'alpha': w['alpha'],
            'durations': durations, 'scaling': w['scaling']}


def load_peaks():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            p = json.load(f)
        return {'hbm': p['hbm_gbs'], 'bf16': p['bf16_tflops'], 'bf16_sustained': p.get('bf16_tflops_sustained'),
                'source': 'measured (MEASURED_PEAKS.json)'}
    except Exception:
        return {'hbm': 6650.0, 'bf16': 1590.0, 'bf16_sustained': 1400.0, 'source': 'fallback (B200_PROFILING.md)'}


def build_model(durations: str, tag: str = 'forward_tacotron'):
    """Random-init weights of the named architecture (there is no checkpoint to download)."""
    from forwardtacotron_b200.configs import default_config
    from forwardtacotron_b200.models import init_tts_model
    torch.manual_seed(SEED)
    model = init_tts_model(default_config(tag))
    if durations == 'spread':           # SURVEY §8d regime B: shift the duration head so durations vary
        with torch.no_grad():
            model.dur_pred.lin.weight.mul_(8.0)
            model.dur_pred.lin.bias.fill_(3.0)
    return model.eval()


def make_tokens(batch, T, rank=0, step=0):
    g = torch.Generator().manual_seed(SEED + rank + 7919 * step)
    return torch.randint(1, 135, (batch, T), generator=g, dtype=torch.int64)


def make_speaker(batch, rank=0):
    g = torch.Generator().manual_seed(SEED + 31 + rank)
    e = torch.randn(batch, 256, generator=g)
    return e / e.norm(dim=1, keepdim=True)            # resemblyzer embeddings are unit-norm


class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region (every ~5 ms through NVML in a thread; falls back
    to an `nvidia-smi -lms` subprocess when the NVML binding is unavailable)."""
    QUERY = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')
    REASON_BITS = {'sw_power_cap': 0x4, 'hw_slowdown': 0x8, 'sw_thermal_slowdown': 0x20, 'hw_thermal_slowdown': 0x40}

    def __init__(self, index):
        self.index, self.sm, self.max_sm, self.reasons = index, [], None, set()
        self.proc = self.thread = self.nvml = self.handle = None
        self.stop = threading.Event()

    def _nvml_handle(self):
        import pynvml
        pynvml.nvmlInit()
        try:
            uuid = str(torch.cuda.get_device_properties(self.index).uuid)
            uuid = uuid if uuid.startswith('GPU-') else 'GPU-' + uuid
            try:
                h = pynvml.nvmlDeviceGetHandleByUUID(uuid)
            except TypeError:
                h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
        return pynvml, h

    def _poll_nvml(self):
        n, h = self.nvml, self.handle
        get_reasons = getattr(n, 'nvmlDeviceGetCurrentClocksEventReasons', None) or \
            getattr(n, 'nvmlDeviceGetCurrentClocksThrottleReasons')
        while not self.stop.is_set():
            try:
                self.sm.append(int(n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM)))
                mask = int(get_reasons(h))
                self.reasons.update(k for k, bit in self.REASON_BITS.items() if mask & bit)
            except Exception:
                pass
            self.stop.wait(0.005)

    def _poll_smi(self):
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for line in self.proc.stdout:
            r = [c.strip() for c in line.split(',')]
            if r and r[0].isdigit():
                self.sm.append(int(r[0]))
            if len(r) > 1 and r[1].isdigit():
                self.max_sm = max(self.max_sm or 0, int(r[1]))
            if len(r) >= 6:
                self.reasons.update(n for n, v in zip(names, r[2:6]) if v.lower() == 'active')

    def __enter__(self):
        try:
            self.nvml, self.handle = self._nvml_handle()
            self.max_sm = int(self.nvml.nvmlDeviceGetMaxClockInfo(self.handle, self.nvml.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._poll_nvml, daemon=True)
            self.thread.start()
            return self
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), f'--query-gpu={self.QUERY}',
                                          '--format=csv,noheader,nounits', '-lms', '20'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._poll_smi, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def __exit__(self, *exc):
        self.stop.set()
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        if self.thread is not None:
            self.thread.join(timeout=2)

    def summary(self):
        sm = sorted(self.sm)
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': self.max_sm,
                'reasons': sorted(self.reasons), 'samples': len(sm),
                'source': 'nvml' if self.nvml is not None else 'nvidia-smi'}


def cpu_forward_timing(sd, batch, runs, warm, threads, wl='config2', device='cpu'):
    """Oracle port of the reference's generate() on the host cores (fused library GRU/LSTM = the kernels the
    reference itself reaches through torch.nn); device='cuda': the same eager PyTorch program on the GPU (cuDNN / cuBLAS
    eager, what `tts_model.to(device)` of gen_forward.py:66-67 gives a user of the reference on this box)."""
    from oracle import ft_oracle as O
    torch.set_num_threads(threads)
    w = WORKLOADS[wl]
    gen = {'forward_tacotron': O.ft_generate, 'fast_pitch': O.fp_generate,
           'multi_forward_tacotron': O.mft_generate}[w['model']]
    x = make_tokens(batch, w['T']).to(device)
    args = (x, make_speaker(batch).to(device)) if w['model'].startswith('multi') else (x,)
    sd = {k: v.to(device) for k, v in sd.items()}
    O.use_library_rnn(True)
    times, frames = [], 0
    try:
        for i in range(warm + runs):
            if device != 'cpu':
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            out = gen(sd, *args, alpha=w['alpha'])
            if device != 'cpu':
                torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            if i >= warm:
                times.append(dt)
            frames = int(out['mel_lens'].sum())
    finally:
        O.use_library_rnn(False)
    return frames, times


def reference_sample(wl, total_runs):
    """Sentences per CPU step: the workload's whole per-GPU batch when `total_runs` of it fit in ~4 minutes of host time
    (config 2: ~1 s per 64x200 step), else a bounded sample of it."""
    w = WORKLOADS[wl]
    est = 1.0 * w['batch'] * w['T'] / (64 * 200)             # ~1 s per 12 800 tokens on 16-32 host cores
    batch = w['batch']
    while batch > 1 and total_runs * est * batch / w['batch'] > 240:
        batch //= 2
    return batch


def run_reference(args, rank, world):
    """`--impl reference` / `--impl eager-gpu`: rank 0 alone runs; the other ranks exit 0 without work."""
    if rank != 0:
        return
    wl = args.workload
    w = WORKLOADS[wl]
    durations = args.durations or w['durations']
    threads = os.cpu_count() or 1
    model = build_model(durations, w['model'])
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    gpu = args.impl == 'eager-gpu'
    batch = w['batch'] if gpu else reference_sample(wl, args.steps + args.warmup)
    frames, times = cpu_forward_timing(sd, batch, args.steps, args.warmup, threads, wl, 'cuda' if gpu else 'cpu')
    sec = sum(times) / len(times)
    value = frames / sec
    sample = f"{batch} of {w['batch']} sentences x {w['T']} phonemes per step (same model, same token distribution)"
    line = {'impl': args.impl, 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': sec * 1e3, 'higher_is_better': True,
            'scaling': w['scaling'], 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': workload_config(wl, durations, world), 'frames_per_step': frames,
            'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': threads, 'kind': 'port', 'sample': sample},
            'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}
    if gpu:
        del line['cpu_baseline']
        line['note'] = ('informative leg: the oracle port of generate() as eager PyTorch on ONE B200 (cuDNN GRU/LSTM, '
                        'cuBLAS/cuDNN convolutions, torch defaults: fp32 matmuls, cuDNN convolutions may use TF32) -- what a user of the reference gets from '
                        'tts_model.to(device) (gen_forward.py:66-67); wall-clock around generate() + synchronize')
    print(json.dumps(line), flush=True)


def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist
    from forwardtacotron_b200 import parallel as par
    from forwardtacotron_b200 import runtime as rt
    dev = torch.device('cuda', local_rank)
    torch.cuda.set_device(dev)
    numa = par.bind_host_to_device(dev) if world > 1 else None
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    peaks = load_peaks()
    wl = args.workload
    w = WORKLOADS[wl]
    durations = args.durations or w['durations']
    T, alpha = w['T'], w['alpha']
    multi = w['model'].startswith('multi')
    model = build_model(durations, w['model']).to(dev)
    handle = rt.Handle.get(dev)
    if w['scaling'] == 'weak':                       # every rank its own batch
        B = w['batch']
        x_pin = make_tokens(B, T, rank).pin_memory()
        semb = make_speaker(B, rank).to(dev) if multi else None
    else:                                            # one global batch, sentence-sharded (parallel.shard_slice)
        sl = par.shard_slice(w['batch'], rank, world)
        B = sl.stop - sl.start
        x_pin = make_tokens(w['batch'], T)[sl].contiguous().pin_memory()
        semb = make_speaker(w['batch'])[sl].contiguous().to(dev) if multi else None
    x_dev = x_pin.to(dev)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2
    # --vary: a different token batch on every step (8 in rotation), so with spread durations the regulated length M
    # changes from call to call like it does with real text (the cached-shape replay is the best case)
    pool = [make_tokens(B, T, rank, step=i).to(dev) for i in range(8)] if args.vary else None
    seen_m, calls = set(), [0]

    def generate(x):
        return model.generate(x, semb, alpha=alpha) if multi else model.generate(x, alpha=alpha)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(args.warmup, 3)):
        out = generate(x_dev)
    torch.cuda.synchronize(dev)
    frames = int((out['dur'] + 0.5).long().sum())                        # valid mel frames per step (this rank)
    if pool is not None:       # every batch of the rotation once (builds / re-points the programs); frames = their mean
        fl = []
        for xb in pool:
            o = generate(xb)
            fl.append(int((o['dur'] + 0.5).long().sum()))
        torch.cuda.synchronize(dev)
        frames = sum(fl) / len(fl)
    M = out['mel'].shape[2]
    host_out = torch.empty(out['mel_post'].shape, dtype=torch.float32).pin_memory()
    b_cap = w['batch'] if w['scaling'] == 'weak' else -(-w['batch'] // world)

    def timed(step_fn):
        evs = []
        barrier()
        for _ in range(args.steps):
            flush.zero_()                                                # evict L2 between timed iterations
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            step_fn()
            b.record()
            evs.append((a, b))
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def step_resident():
        if pool is None:
            generate(x_dev)
        else:
            o = generate(pool[calls[0] % len(pool)])
            calls[0] += 1
            seen_m.add(o['mel'].shape[2])

    exchange = par.LengthExchange(dev) if world > 1 else None
    ship_stream = torch.cuda.Stream(dev)
    counts = [B] * world if w['scaling'] == 'weak' else \
        [(lambda s: s.stop - s.start)(par.shard_slice(w['batch'], r, world)) for r in range(world)]

    trace = [] if os.environ.get('FTB_E2E_TRACE') else None

    def mark():
        if trace is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            trace.append((e, time.perf_counter()))

    def ship(o, shapes):
        """A step's result leaves the GPU: this rank's D2H of mel_post into pinned host memory (PCIe copy engine) and, at
        N > 1, the gather of the variable-length mels to ONE consumer (rank 0) over NVLink, point to point."""
        host_out.copy_(o['mel_post'], non_blocking=True)
        if world > 1:
            par.gather_variable_mels(o['mel_post'], par.mel_lengths(o['dur']), dst=0, shapes=shapes)

    def run_generate(xd):
        if world == 1:
            return generate(xd), None
        # the path's only exchange on the compute side: the regulated lengths cross in one 16-byte NCCL all_gather
        # enqueued from inside generate(); nothing is exchanged on the host and no rank drains its stream
        with exchange:
            o = generate(xd)
        return o, (counts, exchange.lengths())

    def step_e2e_serial():
        """H2D -> generate() -> D2H (+ gather) back to back on one stream: the latency of ONE call."""
        mark()
        o, shapes = run_generate(x_pin.to(dev, non_blocking=True))
        mark()
        ship(o, shapes)
        mark()

    pending = [None]

    def step_e2e():
        """The serving loop: generate() hands out fresh result tensors, so the result of step i-1 is shipped (D2H +
        gather) on a copy stream WHILE step i computes.  The copy stream starts behind this step's opening event and the
        compute stream joins it before the closing one: every timed interval holds one H2D, one generate() and one whole
        D2H (+ gather) -- of the previous step's result; the first interval ships the last warm-up step's."""
        main = torch.cuda.current_stream(dev)
        mark()
        if pending[0] is not None:
            po, pshapes = pending[0]
            ship_stream.wait_stream(main)
            with torch.cuda.stream(ship_stream):
                ship(po, pshapes)
            for t in (po['mel_post'], po['dur']):
                t.record_stream(ship_stream)
        pending[0] = run_generate(x_pin.to(dev, non_blocking=True))
        mark()
        main.wait_stream(ship_stream)
        mark()

    with ClockSampler(local_rank) as clocks:
        l0 = handle.launch_count()
        ms_value = timed(step_resident)
        launches = (handle.launch_count() - l0) // args.steps
        for _ in range(2):                 # untimed: first-use setup of the gather collectives / pinned copy path
            step_e2e_serial()
        ms_e2e_serial = timed(step_e2e_serial)
        for _ in range(2):
            step_e2e()
        if trace is not None:
            del trace[:]
        ms_e2e = timed(step_e2e)
        torch.cuda.synchronize(dev)        # the last step's result drains outside the timed region (see step_e2e)
        if pending[0] is not None:
            with torch.cuda.stream(ship_stream):
                ship_stream.wait_stream(torch.cuda.current_stream(dev))
                ship(*pending[0])
            torch.cuda.synchronize(dev)
            pending[0] = None
        if trace:                       # FTB_E2E_TRACE=1: where an e2e step goes (device and host clocks), per rank
            torch.cuda.synchronize(dev)
            n = 3
            rows = [trace[i:i + n] for i in range(len(trace) - n * args.steps, len(trace), n)]
            dev_ms = [sum(r[j][0].elapsed_time(r[j + 1][0]) for r in rows) / len(rows) for j in range(n - 1)]
            host_ms = [sum((r[j + 1][1] - r[j][1]) * 1e3 for r in rows) / len(rows) for j in range(n - 1)]
            sys.stderr.write(f'e2e trace rank {rank}: device ms per phase {[round(v, 3) for v in dev_ms]} host ms '
                             f'{[round(v, 3) for v in host_ms]} (phases: ship previous result enqueued + h2d + generate, join of the copy stream)\n')
        # per-launch device times of the two programs (buffers still hold the last step's data)
        st = model.__dict__['_accel']
        per = {}
        tok_prog = model.program_for('tok', B, T)
        timed_progs = [('tok', 'tok', tok_prog)]
        if getattr(tok_prog, 'xprog', None) is not None:        # the BiLSTM's x-projection per phoneme (token geometry)
            timed_progs.append(('tok', 'tokx', tok_prog.xprog))
        timed_progs.append(('frame', 'frame', model.program_for('frame', B, M)))
        for side, name, prog in timed_progs:
            for rep in range(4):
                flush.zero_()
                for i, (info, ms) in enumerate(prog.run_timed()):
                    if rep:
                        e = per.setdefault((side, name, i), {'info': info, 'ms': 0.0})
                        e['ms'] += ms / 3.0
        # K5/K6 (length regulator) are launched by generate() itself, outside the two programs: time them with
        # events around the C-ABI calls of three more generate() calls
        lr_ms = {}
        for rep in range(4):
            flush.zero_()
            handle.timing = []
            generate(x_dev)
            torch.cuda.synchronize(dev)
            if rep:
                for name, a, b in handle.timing:
                    lr_ms[name] = lr_ms.get(name, 0.0) + a.elapsed_time(b) / 3.0
            handle.timing = None
    clk = clocks.summary()

    frames_t = torch.tensor([frames], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(frames_t)
    total_frames = float(frames_t.item())
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    kernels = []
    for (side, _name, i), e in per.items():
        info, ms = e['info'], e['ms']
        k = {'side': side, 'label': info['label'], 'kind': info['kind'], 'ms': round(ms, 4)}
        if info['kind'] == 'gemm':
            k['tflops'] = round(info['flops'] / ms / 1e9, 1)
            k['f16_share'] = round((info.get('flops_f16') or 0.0) / max(info['flops'], 1.0), 3)
            k['cta_pairs'] = bool(info.get('cta_pairs'))
        if info['kind'] in ('rnn', 'glue') and info['bytes']:
            k['gbs'] = round(info['bytes'] / ms / 1e6, 1)
        if info['kind'] == 'rnn':
            k['us_per_step'] = round(ms * 1e3 / info['steps'], 3)
            k['handoff'] = info.get('handoff')
            k['kernel_launches'] = info.get('kernel_launches')
        kernels.append((k, info))
    frame_prog = model.program_for('frame', B, M)
    if 'lr' in frame_prog.bufs:
        lr_buf = frame_prog.bufs['lr']
        C = lr_buf.shape[1]
        # bytes the gather really moves: fp32 source rows + durations + the regulated rows in the dtype the product path
        # writes (fp16 operand copy for the BiLSTM's W_ih GEMM: 2 bytes per element, not SURVEY 8d's fp32 figure)
        lr_bytes = {'lr_gather': 4.0 * B * T * C + 4.0 * B * T + float(lr_buf.element_size()) * B * M * C,
                    'lr_scan': 2 * 4.0 * B * T}
    else:
        # per-phoneme x-projection: the regulated sequence is never written.  `lr_gather` is the identity regulator
        # (fp32 prenet rows + projection adds -> fp16 operand rows, T of them per sentence), `lr_index` the frame -> row map
        C = tok_prog.xprog.bufs['h_tok'].shape[1]
        lr_bytes = {'lr_gather': 4.0 * B * T * C + 2.0 * B * T * C, 'lr_scan': 2 * 4.0 * B * T,
                    'lr_index': 4.0 * B * T + 4.0 * B * M}
    for name, ms in lr_ms.items():
        k = {'side': 'lr', 'label': name, 'kind': 'glue', 'ms': round(ms, 4),
             'gbs': round(lr_bytes[name] / ms / 1e6, 1), 'note': 'event-timed around the C-ABI call (includes launch gaps)'}
        kernels.append((k, {'label': name, 'kind': 'glue', 'flops': 0.0, 'bytes': lr_bytes[name]}))
    kernels.sort(key=lambda t: -t[0]['ms'])
    # Roofline of the dominant kernel FUNCTION of the step (all launches of that function summed):
    #   conv_gemm_tc_kernel   -> tensor bound: algorithmic FLOPs (per layer) / summed launch time, against the
    #                            tensor peak of its operand mix: kind::f16 layers against the measured bf16 burst figure,
    #                            kind::tf32 layers against half of it (each launch is timed alone); `peak` is the
    #                            FLOP-weighted harmonic mean, i.e. flops / (time at peak of every layer);
    #   rnn_cluster_kernel / rnn_persistent_kernel (cluster / L2 hand-off of h) -> HBM bound: x-projection
    #                            reads + layer-output writes / summed launch time (the kernels are latency chains;
    #                            their fraction of the HBM peak is reported for completeness).
    groups = {}
    for k, info in kernels:
        fn = {'gemm': 'conv_gemm_tc_kernel',
              'rnn': 'rnn_cluster_kernel' if info.get('handoff') == 'cluster' else 'rnn_persistent_kernel'}.get(k['kind'], k['kind'])
        if k.get('side') == 'lr':
            fn = k['label'] + '_kernel'
        g = groups.setdefault(fn, {'ms': 0.0, 'flops': 0.0, 'flops_f16': 0.0, 'bytes': 0.0, 'launches': 0})
        g['flops_f16'] += info.get('flops_f16') or 0.0
        g['ms'] += k['ms']
        g['flops'] += info.get('flops') or 0.0
        g['bytes'] += info.get('bytes') or 0.0
        g['launches'] += info.get('kernel_launches') or 1
    by_kernel = []
    for fn, g in sorted(groups.items(), key=lambda kv: -kv[1]['ms']):
        if fn == 'conv_gemm_tc_kernel':
            f16, f32 = g['flops_f16'], g['flops'] - g['flops_f16']
            peak = g['flops'] / (f16 / peaks['bf16'] + f32 / (peaks['bf16'] / 2.0))
            ach = g['flops'] / g['ms'] / 1e9
            by_kernel.append({'kernel': fn, 'launches_per_step': g['launches'], 'ms_per_step': round(g['ms'], 4),
                              'bound': 'tensor', 'achieved': ach, 'peak': peak, 'unit': 'TFLOP/s', 'frac': ach / peak,
                              'traffic': None,
                              'f16_flop_share': round(f16 / max(g['flops'], 1.0), 3),
                              'pair_launches_per_step': sum(1 for kk, _ in kernels if kk.get('cta_pairs')),
                              'kernel_note': 'conv_gemm_tc_kernel and its 2-CTA form conv_gemm_tc_pair_kernel '
                                             '(same body, tcgen05.mma.cta_group::2; picked per launch)',
                              'peak_note': f"FLOP-weighted: kind::f16 layers at the {peaks['source']} bf16 burst figure "
                                           f"{peaks['bf16']}, kind::tf32 layers at half of it"})
        elif g['bytes']:
            ach = g['bytes'] / g['ms'] / 1e6
            by_kernel.append({'kernel': fn, 'launches_per_step': g['launches'], 'ms_per_step': round(g['ms'], 4),
                              'bound': 'hbm', 'achieved': ach, 'peak': peaks['hbm'], 'unit': 'GB/s',
                              'frac': ach / peaks['hbm'], 'traffic': None,
                              'peak_note': f"{peaks['source']}; latency-bound when it is a recurrence"})
    # measured DRAM traffic (ncu dram__bytes_read.sum + dram__bytes_write.sum, all launches of the function in one
    # generate(), same workload): committed summary of the launch list under profiles/ (scratch/launch_summary.py)
    try:
        tr_files = sorted(f for f in os.listdir(os.path.join(ROOT, 'profiles')) if f.endswith('_traffic.json'))
        traffic = json.load(open(os.path.join(ROOT, 'profiles', tr_files[-1]))) if tr_files else {}
    except Exception:
        tr_files, traffic = [], {}
    for e in by_kernel:
        t = traffic.get(e['kernel'])
        if t and wl == 'config2' and durations == 'fallback':
            e['traffic'] = t['dram_bytes_per_step']
            e['traffic_note'] = f"DRAM bytes per step over its {t['launches_per_step']} launches, profiles/{tr_files[-1]}"
            e['algorithmic'] = (groups[e['kernel']]['bytes'] if e['bound'] == 'hbm' else None)
    roof = dict(by_kernel[0])
    top = kernels[0][0]
    roof['largest_launch'] = {k: top[k] for k in ('label', 'kind', 'ms') if k in top}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
        nb = reference_sample(wl, 3)
        cf, ct = cpu_forward_timing(sd, nb, runs=2, warm=1, threads=threads, wl=wl)
        cpu = {'value': cf / (sum(ct) / len(ct)), 'unit': UNIT, 'cores': threads, 'kind': 'port',
               'sample': f"{nb} of {w['batch']} sentences x {T} phonemes per step (the --impl reference sample), "
                         'oracle port of generate(), 1 warm-up + 2 timed runs'}

    sec = ms_value / 1e3
    line = {'metric': METRIC, 'value': total_frames * args.steps / sec, 'unit': UNIT, 'n_gpus': world,
            'steps': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': ms_value / args.steps,
            'higher_is_better': True, 'scaling': w['scaling'], 'vs_baseline': None, 'dtype': 'tf32+f16',
            'data': 'synthetic',
            'config': workload_config(wl, durations, world),
            'run': {'frames_per_step_rank0': frames, 'sentences_rank0': B, 'padded_mel_len_rank0': M,
                    'l2': 'flushed between timed iterations (512 MiB memset)',
                    'parallelism': f'sentence-sharded x{world}, weights replicated',
                    'host_cpus_rank0': numa},
            'e2e': {'value': total_frames * args.steps / (ms_e2e / 1e3), 'unit': UNIT,
                    'h2d_bytes_per_step': x_pin.numel() * 8, 'd2h_bytes_per_step': host_out.numel() * 4,
                    'ms_per_step': ms_e2e / args.steps,
                    'pipeline': 'serving loop: the D2H (+ gather to rank 0 at N > 1) of step i-1 runs on a copy stream '
                                'inside the timed interval of step i, beside its compute; every interval holds one H2D, '
                                'one generate() and one whole D2H',
                    'serial_value': total_frames * args.steps / (ms_e2e_serial / 1e3),
                    'serial_ms_per_step': ms_e2e_serial / args.steps,
                    'serial_note': 'H2D -> generate() -> D2H (+ gather) back to back on one stream: one call\'s latency'},
            'inputs': ('one token batch replayed every step' if pool is None else
                       f'8 token batches in rotation (--vary): regulated lengths seen {sorted(seen_m)}'),
            'gpu_launches': int(launches), 'clocks': clk, 'roofline': roof, 'roofline_by_kernel': by_kernel,
            'kernel_ms_sum': round(sum(k['ms'] for k, _ in kernels), 4),
            'kernels': [k for k, _ in kernels]}
    if cpu is not None:
        line['cpu_baseline'] = cpu
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference', 'eager-gpu'])
    ap.add_argument('--workload', default='config2', choices=sorted(WORKLOADS),
                    help='BASELINE.json config: config2 (default, the headline), config3 FastPitch, config4 MultiFT 128 '
                         'sharded, config5 32x4096 sharded')
    ap.add_argument('--durations', default=None, choices=['fallback', 'spread'],
                    help="'fallback': as-is random init -> every duration 2 (forward_tacotron.py:176-177); "
                         "'spread': duration head shifted so durations vary (SURVEY §8d regime B)")
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--vary', action='store_true',
                    help='a different token batch every step (with --durations spread the regulated length M varies)')
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if args.impl in ('reference', 'eager-gpu'):
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == '__main__':
    main()
