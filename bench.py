#!/usr/bin/env python
"""bench.py -- GANtron hot-path benchmark (contract: see the task statement / DESIGN.md section "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Headline metric (BASELINE.json): training mel frames/s of the vanilla GANtron (use_labels=False, use_noise=True)
teacher-forced training step WITH the Discriminator, per-GPU batch 64, synthetic LJSpeech-shaped batch
(T_in = 150 characters, T_out = 800 frames, 80 mels), bf16 operands / fp32 accumulate.

One "step" = one generator iteration (Tacotron2.forward -> Tacotron2Loss incl. guided attention ->
Discriminator.adversarial_loss -> backward -> clip_grad_norm_ -> Adam), i.e. train.py:362-406, followed by one
discriminator iteration on the same batch (real + fake adversarial losses -> backward -> Adam), train.py:301-345.
Frames counted per step = sum(output_lengths) of the generator iteration only.

  value : frames/s with the parsed batch already resident in HBM (CUDA events, max over ranks)
  e2e   : the same through the public API from pinned HOST buffers: model.parse_batch (H2D) ... loss.item() (D2H)
  infer : autoregressive Decoder.inference frames/s at batch 1 and 256 (1000 forced steps), single GPU
  roofline / cpu_baseline / clocks : see DESIGN.md

`--impl reference` times the CPU restatement of the reference path (oracle/, "port") on the box's host cores
on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

B_PER_GPU, T_IN, T_OUT, N_MEL = 64, 150, 800, 80
MFLOP_PER_FRAME_TRAIN = 169.6        # algorithmic FLOP per mel frame of the G step (decoder+postnet x3, D x2)
DEC_STEP_MAC = 23_487_872            # forward MACs of one decoder step per sample
INFER_BYTES_PER_FRAME = 44.85e6      # bf16 decoder weights touched per frame at batch 1


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=float(p["hbm_gbs"]), tf_burst=float(p["bf16_tflops"]), tf_sust=float(p["bf16_tflops_sustained"]),
                    src="measured")
    except Exception:
        return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


class ClockSampler:
    """SM clocks / throttle reasons sampled DURING the timed region (every 100 ms).

    An `nvidia-smi -lms 100` process is started BEFORE the warm-up and its samples are only kept between begin() and stop():
    spawning it at the start of the timed region costs the region the tool's start-up (NVML / driver initialisation, seen
    as a 25 % slower step on a busy host).  Falls back to polling NVML from a thread of this process."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.samples, self.proc, self.nvml = index, [], None, None
        self.active, self.done, self.source = False, False, None

    def start(self):
        # preferred: a separate nvidia-smi process (its NVML polling shares neither the GIL nor this process's driver
        # locks with the launch loop: the in-process NVML thread made the timed region 2-15 % slower than the e2e leg)
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.source = "nvidia-smi"
            threading.Thread(target=self._read, daemon=True).start()
            return
        except Exception:
            self.proc = None
        try:
            import pynvml
            pynvml.nvmlInit()
            # CUDA_VISIBLE_DEVICES remaps CUDA indices; NVML wants the physical device
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = self.index
            if vis:
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                if self.index < len(ids) and ids[self.index].isdigit():
                    phys = int(ids[self.index])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nvml, self.source = pynvml, "nvml"
            threading.Thread(target=self._poll_nvml, daemon=True).start()
        except Exception:
            self.nvml = None

    def begin(self):
        self.samples, self.active = [], True

    def _poll_nvml(self):
        n = self.nvml
        names = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40, "sw_power_cap": 0x4}
        while not self.done:
            if self.active:
                try:
                    sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
                    mx = n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)
                    try:
                        r = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
                    except Exception:
                        r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                    self.samples.append((float(sm), float(mx), [k for k, bit in names.items() if r & bit]))
                except Exception:
                    pass
            time.sleep(0.1)

    def _read(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.proc.stdout:
            if not self.active:
                continue
            parts = [x.strip() for x in line.strip().split(",")]
            if len(parts) < 6:
                continue
            try:
                self.samples.append((float(parts[0]), float(parts[1]),
                                     [n for n, v in zip(names, parts[2:6]) if v.lower().startswith("active")]))
            except ValueError:
                continue

    def stop(self):
        self.active, self.done = False, True
        if self.proc is not None:
            self.proc.terminate()
        if self.source is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi / NVML unavailable"]}
        sm = sorted(s[0] for s in self.samples)
        reasons = sorted({r for s in self.samples for r in s[2]})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.samples[-1][1] if self.samples else None,
                "reasons": reasons, "samples": len(sm), "source": self.source}


# ------------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the oracle port on host cores (bounded sample)
# ------------------------------------------------------------------------------------------------------
CPU_B, CPU_T_IN, CPU_T_OUT = 4, 150, 800      # BASELINE.json configs[0]: the reference's own CPU-runnable case


class CpuPort:
    """BASELINE.json configs[0] on the host cores: vanilla GANtron (use_labels=False, use_noise=True) teacher-forced forward +
    backward, batch 4, T_in = 150, T_out = 800, fp32, all cores.  The decoder / Postnet / Discriminator / losses are the oracle
    port (oracle/gantron_oracle.py: the reference's algorithm as plain torch CPU ops, pinned against outputs of the unmodified
    reference); embedding + text encoder are the stock-PyTorch classes of gantron_b200.model (the reference's own code for
    them is stock PyTorch too).  One step = generator forward, Tacotron2Loss incl. guided attention, frozen-D adversarial
    loss, backward (train.py:362-397)."""

    SAMPLE = ("BASELINE configs[0] exactly: vanilla GANtron teacher-forced fwd+bwd incl. frozen-D adversarial loss, B=4, T_in=150, "
              "T_out=800, fp32 on all host cores; oracle port (torch CPU ops) + stock-PyTorch text encoder")

    def __init__(self, threads: int, seed: int = 1234):
        from oracle import gantron_oracle as O
        from gantron_b200.hparams import HParams
        from gantron_b200 import model as M
        torch.set_num_threads(threads)
        torch.manual_seed(seed)
        self.O = O
        hp = HParams("use_noise=True,use_labels=False")
        self.G = M.Tacotron2(hp).train()
        Dm = M.Discriminator(hp)
        self.d = O.Dims.from_hparams(hp)
        self.sdd = {k: v.detach().clone() for k, v in Dm.state_dict().items()}
        self.batch = O.synthetic_batch(CPU_B, CPU_T_IN, CPU_T_OUT, seed=seed)
        self.gen = torch.Generator().manual_seed(seed)

    def step(self):
        O, d, G, g = self.O, self.d, self.G, self.gen
        text, in_len, mel, gate, spk, emo, out_len = self.batch
        B, T_in, T_out = text.shape[0], text.shape[1], mel.shape[2]
        keep = lambda *s: (torch.rand(*s, generator=g) > 0.5).float()
        masks = dict(pre1=keep(T_out + 1, B, d.prenet), pre2=keep(T_out + 1, B, d.prenet),
                     att=(torch.rand(T_out, B, d.att_rnn, generator=g) > d.p_att).float(),
                     dec=(torch.rand(T_out, B, d.dec_rnn, generator=g) > d.p_dec).float(),
                     post=[keep(B, c, T_out) for c in (512, 512, 512, 512, 80)])
        dk = [keep(B, c[1], (T_out + 19) // 20) for c in O.disc_layer_dims(d)[:4]]
        G.zero_grad()
        sd = {k: v for k, v in G.named_parameters()}
        sd.update({k: v for k, v in G.named_buffers()})
        t0 = time.perf_counter()
        enc = G.encoder(G.embedding(text).transpose(1, 2), in_len)                     # stock PyTorch (out of scope)
        noise = torch.rand(B, 1, d.noise, generator=g).repeat_interleave(T_in, 1)
        memory = torch.cat((enc, noise), -1)
        mo, mpo, go, al = O.generator_from_memory(sd, memory, mel, in_len, out_len, d, masks, training=True)
        ml, gl = O.tacotron2_losses(mo, mpo, go, mel, gate)
        adv = O.adversarial_loss(self.sdd, mpo, out_len, d, dk)
        (ml + gl + adv + 10 * O.attention_loss(al, in_len, out_len)).backward()
        return time.perf_counter() - t0, int(out_len.sum())


def cpu_baseline_median(threads: int, warmup: int = 1, reps: int = 3):
    port = CpuPort(threads)
    for _ in range(warmup):
        port.step()
    ts, frames = [], 0
    for _ in range(reps):
        dt, frames = port.step()
        ts.append(dt)
    ts.sort()
    return frames / ts[len(ts) // 2], ts, frames


def run_reference_arm(args):
    """`--impl reference`: the reference's CPU path for this workload = BASELINE configs[0] (see CpuPort) on all host cores;
    every step is one full configs[0] pass (3,200 frames)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    port = CpuPort(cores)
    for _ in range(max(args.warmup, 0)):
        port.step()
    times, frames = [], 0
    for _ in range(max(1, args.steps)):
        dt, frames = port.step()
        times.append(dt)
    tot = sum(times)
    val = frames * len(times) / tot
    cfg = _config(args.gpus)
    cfg["sample"] = CpuPort.SAMPLE
    cfg["ran"] = {"per_gpu_batch": CPU_B, "t_in": CPU_T_IN, "t_out": CPU_T_OUT, "device": "host CPU", "dtype": "f32"}
    line = {"impl": "reference", "metric": "train_mel_frames_per_s", "value": val, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": len(times), "warmup": max(args.warmup, 0), "ms_per_step": 1e3 * tot / len(times),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg,
            "cpu_baseline": {"value": val, "unit": "frames/s", "cores": cores, "kind": "port", "sample": CpuPort.SAMPLE},
            "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def _config(n_gpus: int):
    return {"workload": "configs[1]: vanilla GANtron (use_labels=False,use_noise=True) teacher-forced training step with "
                        "Discriminator, per-GPU batch 64, T_in=150, T_out=800, 80 mels, bf16",
            "per_gpu_batch": B_PER_GPU, "global_batch": B_PER_GPU * n_gpus, "t_in": T_IN, "t_out": T_OUT,
            "parallelism": f"dp{n_gpus}", "l2": "per-step working set (saved activations ~4 GB) far exceeds the 126 MB L2",
            "step": "1 generator iteration (train.py:362-406) + 1 discriminator iteration (train.py:301-345)"}


# ------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------
class Trainer:
    """The body of train.py's loop (train.py:291-423) against gantron_b200's drop-in model surface."""

    def __init__(self, distributed: bool, seed: int = 1234, hp_string: str = "use_noise=True,use_labels=False"):
        from gantron_b200.hparams import HParams
        from gantron_b200 import model as M
        from gantron_b200.loss_function import Tacotron2Loss
        torch.manual_seed(seed)
        self.hp = HParams(hp_string + ",b200_precision=bf16")
        self.G = M.Tacotron2(self.hp).cuda()
        self.D = M.Discriminator(self.hp).cuda()
        if distributed:
            from gantron_b200.distributed import apply_gradient_allreduce
            self.G = apply_gradient_allreduce(self.G)
            self.D = apply_gradient_allreduce(self.D)
        # train.py:243-250's optimizers; fused=True is the same update as one multi-tensor kernel per step
        self.g_opt = torch.optim.Adam(self.G.parameters(), lr=self.hp.g_learning_rate, weight_decay=self.hp.weight_decay, fused=True)
        self.d_opt = torch.optim.Adam(self.D.parameters(), lr=self.hp.d_learning_rate, weight_decay=self.hp.weight_decay, fused=True)
        self.criterion = Tacotron2Loss()
        self.G.train(); self.D.train()
        self.real, self.fake = 1.0, -1.0

    def step(self, x, y, marks=None):
        """marks: optional list receiving CUDA events at the segment boundaries of the step (scaling probe)."""
        def mark():
            if marks is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                marks.append(e)
        hp, G, D = self.hp, self.G, self.D
        mark()
        # ---- generator iteration (train.py:362-406)
        for p in D.parameters():
            p.requires_grad = False
        G.zero_grad()
        y_pred = G(x)
        generated_mel, lengths = y_pred[1], x[-1]
        mel_loss, gate_loss, attn_loss = self.criterion(y_pred, y, x[1], x[-1])
        adv = self.real * D.adversarial_loss(generated_mel, lengths)
        total = mel_loss + gate_loss + adv + 10 * attn_loss
        mark()
        total.backward()
        mark()
        torch.nn.utils.clip_grad_norm_(G.parameters(), hp.grad_clip_thresh)
        self.g_opt.step()
        mark()
        # ---- discriminator iteration (train.py:301-345)
        for p in D.parameters():
            p.requires_grad = True
        D.zero_grad()
        real_loss = self.real * D.adversarial_loss(y[0], lengths)
        fake_loss = self.fake * D.adversarial_loss(generated_mel.detach(), lengths)
        d_loss = (real_loss + fake_loss) / 2
        torch.nn.utils.clip_grad_norm_(D.parameters(), hp.clipping_value)     # before backward, as train.py:326-327
        d_loss.backward()
        self.d_opt.step()
        mark()
        return total.detach()


def _pinned_batch(B, seed):
    from gantron_b200.synth import synthetic_batch
    return tuple(t.pin_memory() for t in synthetic_batch(B, T_IN, T_OUT, seed=seed))


def bench_combined(steps: int = 3):
    """BASELINE configs[3]: combined labels + noise GANtron (use_labels=True, use_noise=True, VESUS-style speaker embedding and
    5-emotion conditioning; D_mem = 1093), training step with Discriminator, batch 64, bf16, one GPU."""
    tr = Trainer(False, hp_string="use_noise=True,use_labels=True,vesus_path=synthetic")
    batch = _pinned_batch(B_PER_GPU, seed=4321)
    x, y = tr.G.parse_batch(batch)
    for _ in range(3):
        tr.step(x, y)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = tr.step(x, y)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return {"workload": "configs[3]: combined labels+noise GANtron training step with Discriminator, batch 64, T_in=150, T_out=800, bf16",
            "value": int(batch[-1].sum()) / (ms * 1e-3), "unit": "frames/s", "ms_per_step": ms, "steps": steps, "warmup": 3,
            "last_loss": float(loss)}


def bench_inference(steps_frames: int = 1000, pk=None):
    """BASELINE configs[2]: autoregressive inference, max_decoder_steps = 1000 (gate forced not to fire: random weights never
    stop), T_in = 150, batch 1 and batch 256, through `Decoder.inference` (the module API `Tacotron2.inference`
    calls, model.py:714).  `value`: encoder outputs resident in HBM; `e2e`: encoder outputs + style in pinned host memory,
    H2D copy, decode, D2H copy of the mel spectrogram, wall clock."""
    from gantron_b200.hparams import HParams
    from gantron_b200.model import Decoder
    torch.manual_seed(0)
    hp = HParams(f"use_noise=True,use_labels=False,b200_precision=bf16,max_decoder_steps={steps_frames},gate_threshold=2.0")
    dec = Decoder(hp).cuda().eval()
    pk = pk or _peaks()
    out = {}
    for Bi in (1, 256):
        enc_h = (torch.randn(Bi, T_IN, 512) * 0.5).pin_memory()
        sty_h = torch.rand(Bi, 1, 512).repeat_interleave(T_IN, 1).contiguous().pin_memory()
        enc, sty = enc_h.cuda(), sty_h.cuda()
        mel_host = torch.empty(Bi, N_MEL, steps_frames, dtype=torch.float32).pin_memory()   # D2H destination of the e2e leg
        best = best_e2e = None
        for it in range(4):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            with torch.no_grad():
                mel, gate, align = dec.inference(enc, sty)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            t0 = time.perf_counter()
            with torch.no_grad():
                mel2, _, _ = dec.inference(enc_h.cuda(non_blocking=True), sty_h.cuda(non_blocking=True))
                assert mel2.shape == mel_host.shape
                mel_host.copy_(mel2, non_blocking=True)     # into pinned host memory (a pageable .cpu() copy costs 4x the PCIe time)
                torch.cuda.synchronize()
            wall = (time.perf_counter() - t0) * 1e3
            assert mel.shape[-1] == steps_frames and float(mel_host[0, 0, -1]) == float(mel_host[0, 0, -1])
            if it > 0:
                best = ms if best is None else min(best, ms)
                best_e2e = wall if best_e2e is None else min(best_e2e, wall)
        us = 1e3 * best / steps_frames
        r = {"frames_per_s": Bi * steps_frames / (best * 1e-3), "us_per_frame_step": us, "frames": steps_frames,
             "api": "Decoder.inference",
             "e2e": {"frames_per_s": Bi * steps_frames / (best_e2e * 1e-3), "us_per_frame_step": 1e3 * best_e2e / steps_frames,
                     "h2d_bytes": int(enc_h.numel() * 4 + sty_h.numel() * 4), "d2h_bytes": int(Bi * N_MEL * steps_frames * 4)}}
        if Bi == 1:
            # algorithmic bytes per frame: the bf16 decoder weights touched by one step (SURVEY 8d, 44.85 MB) against HBM
            # streaming.  The batch-1 kernel keeps them on chip (tensor memory), so this is a fraction of a roofline it does
            # not sit on; the working bound is 3 grid barriers + the on-chip mat-vecs per frame (profiles/r2_infer1_phases.md)
            ach = INFER_BYTES_PER_FRAME / (us * 1e-6) / 1e9
            r["roofline"] = {"bound": "hbm", "achieved": ach, "peak": pk["hbm"], "unit": "GB/s", "frac": ach / pk["hbm"],
                             "kernel": "decoder_infer1_kernel", "note": "weights resident in tensor memory: latency-bound, target < 20 us/frame"}
        else:
            # max(12.0 GFLOP / sustained bf16 peak, (44.85 MB weights + 49 MB memory / processed memory) / HBM)
            t_tensor = Bi * DEC_STEP_MAC * 2.0 / (pk["tf_sust"] * 1e12)
            t_hbm = (INFER_BYTES_PER_FRAME + Bi * T_IN * (512 + 128) * 2.0) / (pk["hbm"] * 1e9)
            floor_us = max(t_tensor, t_hbm) * 1e6
            r["roofline"] = {"bound": "hbm" if t_hbm > t_tensor else "tensor", "floor_us_per_step": floor_us,
                             "achieved_us_per_step": us, "frac": floor_us / us,
                             "achieved": Bi * DEC_STEP_MAC * 2.0 / (us * 1e-6) / 1e12, "peak": pk["tf_sust"], "unit": "TFLOP/s",
                             "kernel": "decoder_fwd2_kernel<256, FAST>"}
        out[f"b{Bi}"] = r
    return out


def eager_port_on_gpu(steps: int = 2):
    """The practical bar SURVEY 8d asks for, as far as it can travel: the reference algorithm as EAGER PyTorch on this B200 at
    the headline shape (B = 64, T_in = 150, T_out = 800, fp32, cuBLAS / ATen kernels, ~40 launches per decoder step).  The
    reference's own source cannot be copied to the GPU box, so this runs the oracle PORT's torch ops on CUDA tensors (labelled
    `port`): generator from encoder outputs -> losses -> frozen-D adversarial loss -> backward."""
    from oracle import gantron_oracle as O
    from gantron_b200.hparams import HParams
    from gantron_b200 import model as M
    torch.manual_seed(1234)
    hp = HParams("use_noise=True,use_labels=False")
    dev = torch.device("cuda")
    G, Dm = M.Tacotron2(hp), M.Discriminator(hp)
    d = O.Dims.from_hparams(hp)
    sd = {k: v.detach().to(dev).requires_grad_(v.dtype.is_floating_point) for k, v in G.state_dict().items()}
    sdd = {k: v.detach().to(dev) for k, v in Dm.state_dict().items()}
    B, T_in, T_out = B_PER_GPU, T_IN, T_OUT
    text, in_len, mel, gate, spk, emo, out_len = O.synthetic_batch(B, T_in, T_out, seed=1234)
    mel, gate = mel.to(dev), gate.to(dev)
    g = torch.Generator(device=dev).manual_seed(1)
    keep = lambda *s: (torch.rand(*s, generator=g, device=dev) > 0.5).float()
    times = []
    for it in range(steps + 1):
        memory = torch.cat((torch.randn(B, T_in, 512, generator=g, device=dev) * 0.5,
                            torch.rand(B, 1, d.noise, generator=g, device=dev).repeat_interleave(T_in, 1)), -1).requires_grad_(True)
        masks = dict(pre1=keep(T_out + 1, B, d.prenet), pre2=keep(T_out + 1, B, d.prenet),
                     att=(torch.rand(T_out, B, d.att_rnn, generator=g, device=dev) > d.p_att).float(),
                     dec=(torch.rand(T_out, B, d.dec_rnn, generator=g, device=dev) > d.p_dec).float(),
                     post=[keep(B, c, T_out) for c in (512, 512, 512, 512, 80)])
        dk = [keep(B, c[1], (T_out + 19) // 20) for c in O.disc_layer_dims(d)[:4]]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        mo, mpo, go, al = O.generator_from_memory(sd, memory, mel, in_len.to(dev), out_len.to(dev), d, masks, training=True)
        ml, gl = O.tacotron2_losses(mo, mpo, go, mel, gate)
        adv = O.adversarial_loss(sdd, mpo, out_len, d, dk)
        (ml + gl + adv).backward()
        torch.cuda.synchronize()
        if it:
            times.append(time.perf_counter() - t0)
        for v in sd.values():
            v.grad = None
    ms = 1e3 * sum(times) / len(times)
    return {"value": B * T_out / (ms * 1e-3), "unit": "frames/s", "ms_per_step": ms, "kind": "port", "dtype": "f32",
            "sample": f"oracle port's torch ops on CUDA tensors (eager, cuBLAS/ATen), generator fwd+bwd from encoder outputs incl. losses "
                      f"and frozen-D adversarial loss, B={B}, T_in={T_in}, T_out={T_out}, mean of {len(times)} after 1 warm-up; no D "
                      f"iteration, no optimizer, no text encoder"}


def run_ours(args):
    import torch.distributed as dist
    from gantron_b200 import _lib
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (impl ours) needs a CUDA device: gantron_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    distributed = world > 1
    if distributed:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        import datetime
        # short watchdog: a collective mismatch must fail in two minutes, not hang the box for ten
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=120))
    _lib.load()
    tr = Trainer(distributed)
    # weak scaling by default (per-GPU batch 64, the configuration the metric is quoted on); --global-batch G runs BASELINE
    # configs[4] as written (global batch 512 => per-GPU 512 / N: strong scaling)
    b_gpu = B_PER_GPU if not args.global_batch else max(1, args.global_batch // world)
    batch = _pinned_batch(b_gpu, seed=1234 + rank)
    frames = int(batch[-1].sum())
    h2d = sum(t.numel() * t.element_size() for t in batch)

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(k, e2e):
        x = y = None
        if not e2e:
            x, y = tr.G.parse_batch(batch)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        last = None
        for _ in range(k):
            if e2e:
                x, y = tr.G.parse_batch(batch)
            loss = tr.step(x, y)
            if e2e:
                last = loss.item()
        e1.record()
        barrier()
        wall = time.perf_counter() - t0
        ms = e0.elapsed_time(e1)
        if e2e:
            ms = max(ms, wall * 1e3)        # e2e includes host work: take the larger of device span and wall clock
        t = torch.tensor([ms], device="cuda")
        if distributed:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), last

    # warm-up (>= 3): also builds the weight packs / allocator pools
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    timed(max(args.warmup, 3), False)
    if rank == 0:
        sampler.begin()
    _lib.reset_launch_count()
    ms_total, _ = timed(args.steps, False)
    launches = _lib.launch_count()
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e, last_loss = timed(args.steps, True)

    probe = None
    if args.scale_probe:
        # per-rank segment times of the step (forward+losses | G backward incl. gradient all-reduce | clip + Adam | D step)
        x, y = tr.G.parse_batch(batch)
        seg = [[] for _ in range(4)]
        for _ in range(6):
            marks = []
            barrier()
            tr.step(x, y, marks)
            torch.cuda.synchronize()
            for i in range(4):
                seg[i].append(marks[i].elapsed_time(marks[i + 1]))
        mine = torch.tensor([sum(s_[1:]) / len(s_[1:]) for s_ in seg], device="cuda")
        allr = [torch.zeros_like(mine) for _ in range(world)]
        if distributed:
            dist.all_gather(allr, mine)
        else:
            allr = [mine]
        probe = {"segments": ["G forward + losses", "G backward (+ gradient all-reduce)", "clip + Adam (G)", "D step"],
                 "ms_per_rank": [[round(float(v), 3) for v in r_] for r_ in allr]}
    value = frames * world * args.steps / (ms_total * 1e-3)
    e2e_val = frames * world * args.steps / (ms_e2e * 1e-3)
    line = None
    pk = _peaks()
    # dominant kernels, measured live inside full steps; EVERY rank runs it (the steps contain the gradient all-reduce)
    roof = kernel_roofline(tr, batch, pk, (clocks or {}).get("sm_mhz") if rank == 0 else None)
    if rank == 0:
        cfg = _config(world)
        if args.global_batch:
            cfg.update(per_gpu_batch=b_gpu, global_batch=b_gpu * world,
                       workload=f"configs[4]: data-parallel GANtron training, global batch {b_gpu * world} (per-GPU {b_gpu}), bf16")
        line = {"metric": "train_mel_frames_per_s", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
                "scaling": "strong" if args.global_batch else "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": cfg,
                "e2e": {"value": e2e_val, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                        "ms_per_step": ms_e2e / args.steps, "last_loss": last_loss},
                "gpu_launches": int(launches), "clocks": clocks, "roofline": roof,
                "step_roofline": {"bound": "tensor", "achieved": MFLOP_PER_FRAME_TRAIN * 1e6 * (value / world) / 1e12,
                                  "peak": pk["tf_sust"], "unit": "TFLOP/s",
                                  "frac": MFLOP_PER_FRAME_TRAIN * 1e6 * (value / world) / 1e12 / pk["tf_sust"],
                                  "peak_source": pk["src"], "note": "whole G+D step, algorithmic 169.6 MFLOP/frame"}}
    if rank == 0 and probe is not None:
        line["scale_probe"] = probe
    if world == 1 and rank == 0:
        del tr
        torch.cuda.empty_cache()
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):      # (`Warning! Reached max decoder steps`, model.py:489, is not JSON)
            if not args.no_infer:
                line["infer"] = bench_inference(pk=pk)
            if not args.no_extra:
                line["combined_config"] = bench_combined()
                torch.cuda.empty_cache()
                line["eager_port_on_b200"] = eager_port_on_gpu()
                torch.cuda.empty_cache()
        if not args.no_cpu:
            cores = os.cpu_count() or 1
            val, ts, fr = cpu_baseline_median(cores)
            line["cpu_baseline"] = {"value": val, "unit": "frames/s", "cores": cores, "kind": "port",
                                    "sample": CpuPort.SAMPLE + f"; 1 warm-up + median of {len(ts)} passes ({fr} frames each)",
                                    "seconds_per_pass": ts}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if distributed:
        dist.barrier()
        dist.destroy_process_group()


# Algorithmic work of the BPTT kernel per sample-frame (SURVEY 8d shapes, vanilla dims): data-gradient GEMMs
# W_dec^T (2560 x 4096) + W_att^T (1536 x 4096), Wq^T (1024 x 128), location filter backward (d[prev|cum] and dU: 2 x 912,000),
# energies / context backward (2 x 172,800) -- MACs; FLOP = 2 x MAC.  Prenet / constant-column rows are deferred to batched GEMMs.
BWD_KERNEL_MAC = (2560 + 1536) * 4096 + 131_072 + 2 * 912_000 + 2 * 172_800
FWD_KERNEL_MAC = DEC_STEP_MAC - 86_016 - 165_888 - 4096 * 256 - 4096 * 1024 - 0   # minus prenet, projection, hoisted prenet/constant columns
# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` captures (profiles/r2_ncu_decoder.md)
NCU_TRAFFIC = {"decoder_fwd3_kernel": 6.92e9, "decoder_bwd2_kernel": 7.51e9}   # B = 64, T_in = 150, T_out = 800
# cycles-per-frame floors of the two recurrent kernels: L2 -> SM stream of the operands that cannot stay
# on chip at the measured ~30 B/clk/SM plus two grid barriers of ~2k cycles
FLOOR_CYCLES_PER_FRAME = {"decoder_fwd3_kernel": 11_000, "decoder_bwd2_kernel": 13_000}


def kernel_roofline(tr, batch, pk, sm_mhz=None):
    """Live CUDA-event timing (on the launching stream) of every launch of the two persistent decoder kernels inside full
    training steps; `roofline` describes the dominant one (the BPTT kernel).  Both are LATENCY-bound (800 dependent frames x 2
    grid barriers, M = batch = 64 per frame): next to the tensor-pipe fraction the line carries cycles per frame against the
    kernel's own floor."""
    from gantron_b200 import decoder_ops, ops
    x, y = tr.G.parse_batch(batch)
    decoder_ops.KERNEL_EVENTS = {}
    for _ in range(3):
        tr.step(x, y)
    torch.cuda.synchronize()
    ev = decoder_ops.KERNEL_EVENTS
    decoder_ops.KERNEL_EVENTS = None
    frames = int(batch[-1].sum())
    n_frames = int(batch[2].shape[2])
    mhz = sm_mhz or 1965.0
    out = {}
    for name, mac in (("decoder_bwd2_kernel", BWD_KERNEL_MAC), ("decoder_fwd3_kernel", FWD_KERNEL_MAC)):
        ms = [a.elapsed_time(b) for a, b in ev.get(name, [])]
        if not ms:
            continue
        avg = sum(ms) / len(ms)
        ach = 2.0 * mac * frames / (avg * 1e-3) / 1e12
        cyc = avg * 1e-3 / n_frames * mhz * 1e6
        out[name] = {"bound": "latency", "kernel": name, "achieved": ach, "peak": pk["tf_sust"], "unit": "TFLOP/s",
                     "frac": ach / pk["tf_sust"], "traffic": NCU_TRAFFIC.get(name), "peak_source": pk["src"],
                     "launches_timed": len(ms), "ms_per_launch": avg, "us_per_frame": 1e3 * avg / n_frames,
                     "cycles_per_frame": cyc, "floor_cycles_per_frame": FLOOR_CYCLES_PER_FRAME[name],
                     "latency_frac": FLOOR_CYCLES_PER_FRAME[name] / cyc, "sm_mhz": mhz,
                     "algorithmic_flop_per_launch": 2.0 * mac * frames,
                     "note": "recurrent kernel: `achieved`/`frac` are the tensor-pipe view (algorithmic FLOP / sustained bf16 peak); the "
                             "bound that applies is the dependency chain of the frames, see cycles_per_frame vs floor and "
                             "profiles/r2_phase_cycles.md"}
    # the batched (time-independent) GEMMs are the ones the >= 70 % tensor-pipe target is about: time the largest one live
    TB = B_PER_GPU * T_OUT
    dy = torch.randn(TB, 4096, device="cuda").to(torch.bfloat16)
    xx = torch.randn(TB, 1024, device="cuda").to(torch.bfloat16)
    o = torch.empty(4096, 1024, device="cuda")
    best = None
    for it in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.gemm_tn(dy, xx, o, block_n=256)
        e1.record()
        torch.cuda.synchronize()
        if it:
            best = e0.elapsed_time(e1) if best is None else min(best, e0.elapsed_time(e1))
    tf = 2.0 * 4096 * 1024 * TB / (best * 1e-3) / 1e12
    assert out, "kernel timing hook recorded no launch of the persistent decoder kernels"
    roof = out.get("decoder_bwd2_kernel") or next(iter(out.values()))
    roof["other_kernels"] = {k: v for k, v in out.items() if k != roof["kernel"]}
    roof["batched_gemm"] = {"kernel": "gemm_tc_kernel<256> weight gradient M=4096 N=1024 K=51200 (MN-major bf16)", "bound": "tensor",
                            "achieved": tf, "peak": pk["tf_burst"], "unit": "TFLOP/s", "frac": tf / pk["tf_burst"], "ms": best}
    return roof


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-infer", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the configs[3] and eager-PyTorch-on-B200 legs")
    ap.add_argument("--scale-probe", action="store_true", help="add per-rank segment times of the step (profiles/r2_scale_breakdown.md)")
    ap.add_argument("--global-batch", type=int, default=0, help="configs[4] as written: global batch split over the GPUs (strong scaling)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        # convenience: relaunch under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29517"), __file__,
               "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup)]
        if args.global_batch:
            cmd += ["--global-batch", str(args.global_batch)]
        for flag, on in (("--scale-probe", args.scale_probe), ("--no-infer", args.no_infer), ("--no-cpu", args.no_cpu),
                         ("--no-extra", args.no_extra)):
            if on:
                cmd.append(flag)
        sys.exit(subprocess.call(cmd))
    run_ours(args)


if __name__ == "__main__":
    main()
