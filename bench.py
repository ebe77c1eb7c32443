#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (BASELINE.json):

    ResNet-50 training step (all layers trainable, SGD lr 0.001, NLL loss -- resnet.py:17-25 recipe on the torchvision
    graph the reference uses at pytorch_training_inference_on_image.ipynb:389), bf16 tensor-core mode,
    batch 256 per GPU, synthetic Imagenette-shaped input 3x160x160 (configs[1]); weak scaling under
    b200net.parallel.DistributedDataParallel for --gpus > 1 (configs[2]).

    python bench.py --gpus N --steps K --warmup W            # our arm (one JSON line from rank 0)
    python bench.py --impl reference --gpus N ...            # the reference's CPU path on the host cores

value  : images/s, whole job, inputs resident in HBM when the timed region starts (CUDA events, max over ranks)
e2e    : the reference's loop verbatim -- inputs.to(device) from pinned host memory and loss.item() every step
roofline: the tcgen05 conv/GEMM kernel family, algorithmic FLOPs (12.398 GFLOP/img) over the
         summed per-launch CUDA-event durations of that kernel, against the measured sustained bf16 peak
cpu_baseline: the oracle (stock torch CPU fp32, the reference's statements) on a bounded sample, rank 0, N=1 only
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "resnet50_train_images_per_sec"
UNIT = "img/s"
PER_GPU_BATCH = 256
HW = 160
GFLOP_PER_IMG_TRAIN = 12.398     # BASELINE.md section 3 / SURVEY 8(d): fwd + dgrad + wgrad, no stem dgrad
# roofline.traffic: DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum, `ncu --set full`) of ONE NAMED launch of
# the dominant kernel -- the forward 64->256 1x1 convolution of layer1 at 40x40 (class c of SURVEY 8(a'), four per
# step and the longest conv launch of the step) -- next to the ALGORITHMIC bytes of that same launch (read x, write y,
# read w once).  Source: profiles/r02_ncu_named_launches.txt.
CONV_TRAFFIC_LAUNCH = "conv_sm100_kernel<256,3,fprop,1,2> with fused BatchNorm statistics: 256 x 40x40, 64->256, 1x1 (class c)"
CONV_TRAFFIC_BYTES_PER_LAUNCH = None            # filled from profiles/r02_ncu_named_launches.json when present
CONV_ALGO_BYTES_PER_LAUNCH = 2.0 * (256 * 40 * 40 * 64 + 256 * 40 * 40 * 256 + 256 * 64)
BERT_GEMM_TRAFFIC_BYTES_PER_LAUNCH = 36.8e6


# ResNet-50 convolutions at 160 x 160 by layer class (SURVEY.md 8(a')): spatial size, Cin, Cout, kernel, stride, pad, count
RESNET50_CONV_CLASSES = [
    (40, 64, 64, 1, 1, 0, 1), (40, 64, 64, 3, 1, 1, 3), (40, 64, 256, 1, 1, 0, 4), (40, 256, 64, 1, 1, 0, 2),
    (40, 256, 128, 1, 1, 0, 1), (40, 128, 128, 3, 2, 1, 1), (20, 128, 512, 1, 1, 0, 4), (40, 256, 512, 1, 2, 0, 1),
    (20, 512, 128, 1, 1, 0, 3), (20, 128, 128, 3, 1, 1, 3), (20, 512, 256, 1, 1, 0, 1), (20, 256, 256, 3, 2, 1, 1),
    (10, 256, 1024, 1, 1, 0, 6), (20, 512, 1024, 1, 2, 0, 1), (10, 1024, 256, 1, 1, 0, 5), (10, 256, 256, 3, 1, 1, 5),
    (10, 1024, 512, 1, 1, 0, 1), (10, 512, 512, 3, 2, 1, 1), (5, 512, 2048, 1, 1, 0, 3), (10, 1024, 2048, 1, 2, 0, 1),
    (5, 2048, 512, 1, 1, 0, 2), (5, 512, 512, 3, 1, 1, 2),
]


def conv_per_layer_roofline_ms(batch, tflops, hbm_gbs):
    """Sum over the 53 convolutions x {fprop, dgrad, wgrad} of max(tensor time, HBM time) of that launch: flops at the
    given bf16 peak, algorithmic bytes (activations in and out in bf16, weights once; fp32 weight gradient) at the given
    HBM rate.  Most 1x1 layers are HBM-bound, the 3x3 ones tensor-bound: a single-bound roofline misstates both."""
    peak, hbm = tflops * 1e12, hbm_gbs * 1e9
    total = 0.0
    for H, C, K, R, st, pad, cnt in RESNET50_CONV_CLASSES:
        Ho = (H + 2 * pad - R) // st + 1
        x, y, w = batch * H * H * C, batch * Ho * Ho * K, K * C * R * R
        flops = 2.0 * batch * Ho * Ho * K * C * R * R
        t_f = max(flops / peak, 2.0 * (x + y + w) / hbm)
        t_w = max(flops / peak, (2.0 * (x + y) + 4.0 * w) / hbm)
        total += cnt * (2 * t_f + t_w)
    # stem 3 -> 64, 7x7 / 2 (no dgrad: the image needs no gradient); input counted as 8-byte bf16 pixels
    x, y, w = batch * 160 * 160 * 4, batch * 80 * 80 * 64, 64 * 3 * 49
    flops = 2.0 * batch * 80 * 80 * 64 * 147
    total += max(flops / peak, 2.0 * (x + y + w) / hbm) + max(flops / peak, (2.0 * (x + y) + 4.0 * w) / hbm)
    return total * 1e3


def _named_launch_traffic():
    p = os.path.join(ROOT, "profiles", "r02_ncu_named_launches.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("conv_c_fprop", {}).get("dram_bytes")
        except Exception:
            return None
    return None
CPU_SAMPLE_BATCH = 32            # configs[0] batch: the reference's own CPU-runnable case
WORKLOAD = ("resnet50 train step (Mode A: all layers trainable, SGD 0.001, NLL), batch %d/GPU, 3x%dx%d "
            "(BASELINE.json configs[1]; DDP = configs[2] when n_gpus > 1)")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"tflops": d.get("bf16_tflops_sustained", d.get("bf16_tflops")), "hbm": d.get("hbm_gbs"), "src": "measured"}
    return {"tflops": 1400.0, "hbm": 6650.0, "src": "fallback"}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {
                getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
                getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
                getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
                getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
                getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
            }
            while not self._stop_evt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(0.05)
        except Exception as e:  # NVML unavailable: report that instead of inventing numbers
            self.reasons.add("nvml_unavailable:%s" % type(e).__name__)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


def cpu_reference_run(steps, warmup, batch=CPU_SAMPLE_BATCH):
    """The reference's training-step statements on stock torch CPU fp32, all host threads."""
    import torch
    from oracle import torch_oracle as O
    cores = O.cpu_threads()
    torch.set_num_threads(cores)
    model = O.build_reference_model(dropout=True)
    opt, crit = O.mode_a(model)
    model.train()
    x, y = O.make_batch(batch, HW, 0)
    for _ in range(warmup):
        O.train_step(model, opt, crit, x, y)
    t0 = time.perf_counter()
    for _ in range(steps):
        O.train_step(model, opt, crit, x, y)
    dt = time.perf_counter() - t0
    return batch * steps / dt, dt / steps * 1e3, cores


def run_reference(args, rank):
    if rank != 0:
        return
    ips, ms, cores = cpu_reference_run(args.steps, args.warmup)
    sample = ("stock torch %s CPU fp32 executing the reference train-step statements "
              "(pytorch_training_inference_on_image.ipynb:667-673, all layers trainable, SGD 0.001); each step = "
              "%d images of the %dx3x%dx%d workload" % (__import__("torch").__version__, CPU_SAMPLE_BATCH,
                                                         PER_GPU_BATCH, HW, HW))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": ips, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD % (PER_GPU_BATCH, HW, HW), "cpu_sample_batch": CPU_SAMPLE_BATCH},
        "cpu_baseline": {"value": ips, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": ips, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


# ------------------------------------------------------------------ BERT / IMDB leg (BASELINE.json configs[4], C5)
BERT_BATCH, BERT_SEQ = 32, 256          # per GPU (pytorch_on_language_distr.py:134 batch_size=32; SURVEY C5 S=256)
BERT_GFLOP_PER_SEQ_TRAIN = 137.7        # 3 x 12 x (2 x 7 077 888 + 4 x S x 768) x S at S=256
BERT_WORKLOAD = ("bert-base-uncased sequence classifier train step (pytorch_on_language_distr.py:257-275: forward with "
                 "labels, backward, clip_grad_norm_ 1.0, AdamW 2e-5/1e-8, linear schedule), batch %d/GPU, seq %d, "
                 "dropout 0.1 (BASELINE.json configs[4])")


def bert_cpu_reference_run(steps, warmup, batch, seq=BERT_SEQ):
    """The reference's BERT train-step statements on transformers + stock torch CPU fp32, all host threads."""
    import torch
    from oracle import bert_oracle as BO
    from oracle.torch_oracle import cpu_threads
    cores = cpu_threads()
    torch.set_num_threads(cores)
    model = BO.build_reference_model(dropout=0.1)
    opt, sched = BO.make_optimizer(model, 1000)
    ids, mask, labels = BO.make_batch(batch, seq, 0)
    model.train()
    for _ in range(warmup):
        BO.train_step(model, opt, sched, ids, mask, labels)
    t0 = time.perf_counter()
    for _ in range(steps):
        BO.train_step(model, opt, sched, ids, mask, labels)
    dt = time.perf_counter() - t0
    return batch * steps / dt, dt / steps * 1e3, cores


def bert_leg(dev, rank, world, distributed, steps, warmup, timed, with_roofline):
    """One rank's share of the C5 step through the reference's own statements; returns the numbers of the leg."""
    import ctypes
    import torch
    import b200net
    from b200net import bert
    from oracle import bert_oracle as BO     # synthetic batch generator only
    lib = b200net._C.lib
    torch.manual_seed(BO.SEED_WEIGHTS)
    model = bert.BertForSequenceClassification(bert.BertConfig(num_labels=2)).to(dev).train()
    optimizer = bert.AdamW(model.parameters(), lr=2e-5, eps=1e-8)
    scheduler = bert.get_linear_schedule_with_warmup(optimizer, num_warmup_steps=0, num_training_steps=10000)
    net = model
    if distributed:
        net = b200net.parallel.DistributedDataParallel(model, device_ids=[dev.index], output_device=dev.index)
    ids_h, mask_h, lab_h = BO.make_batch(BERT_BATCH, BERT_SEQ, 2000 + rank)
    ids_h, mask_h, lab_h = ids_h.pin_memory(), mask_h.pin_memory(), lab_h.pin_memory()
    ids_d, mask_d, lab_d = ids_h.to(dev), mask_h.to(dev), lab_h.to(dev)

    def body(b_input_ids, b_input_mask, b_labels):
        # pytorch_on_language_distr.py:257-275
        net.zero_grad()
        outputs = net(b_input_ids, token_type_ids=None, attention_mask=b_input_mask, labels=b_labels)
        loss = outputs[0]
        loss.backward()
        b200net.utils.clip_grad_norm_(model.parameters(), 1.0)
        optimizer.step()
        scheduler.step()
        return loss

    def step_resident():
        return body(ids_d, mask_d, lab_d)

    def step_e2e():
        return body(ids_h.to(dev, non_blocking=True), mask_h.to(dev, non_blocking=True),
                    lab_h.to(dev, non_blocking=True)).item()

    step_resident()
    torch.cuda.synchronize()
    l0 = lib.b200_launch_count()
    step_resident()
    launches = lib.b200_launch_count() - l0
    for _ in range(max(warmup, 3) + 3):          # includes the eager calls that precede graph capture
        step_resident()
    ms = timed(step_resident, steps)
    for _ in range(3):
        step_e2e()
    e2e_ms = timed(step_e2e, steps)
    out = {"workload": BERT_WORKLOAD % (BERT_BATCH, BERT_SEQ), "batch_per_gpu": BERT_BATCH, "seq_len": BERT_SEQ,
           "seq_s": world * BERT_BATCH * steps / (ms * 1e-3), "tokens_s": world * BERT_BATCH * BERT_SEQ * steps / (ms * 1e-3),
           "ms_per_step": ms / steps,
           "e2e": {"value": world * BERT_BATCH * steps / (e2e_ms * 1e-3), "unit": "seq/s",
                   "h2d_bytes_per_step": (ids_h.numel() + mask_h.numel() + lab_h.numel()) * 8, "d2h_bytes_per_step": 4,
                   "ms_per_step": e2e_ms / steps},
           "gpu_launches_per_step": int(launches),
           "tensor_util_step": BERT_GFLOP_PER_SEQ_TRAIN * 1e9 * BERT_BATCH * steps / (ms * 1e-3) / 1e12 / peaks()["tflops"]}
    if with_roofline:
        prof_steps = 2
        torch.cuda.synchronize()
        b200net.set_cuda_graphs(False)            # per-launch events need the eager path
        lib.b200_gemm_timing_enable(1)
        for _ in range(prof_steps):
            step_resident()
        torch.cuda.synchronize()
        tms, n, fl = ctypes.c_double(), ctypes.c_int64(), ctypes.c_double()
        lib.b200_gemm_timing_read(ctypes.byref(tms), ctypes.byref(n), ctypes.byref(fl))
        lib.b200_gemm_timing_enable(0)
        b200net.set_cuda_graphs(True)
        pk = peaks()
        achieved = fl.value / (tms.value * 1e-3) / 1e12
        out["roofline"] = {"bound": "tensor", "kernel": "conv_sm100_kernel as flat 1x1 (the encoder's Linear layers: "
                           "fprop / dgrad / wgrad)", "achieved": achieved, "peak": pk["tflops"], "unit": "TFLOP/s",
                           "frac": achieved / pk["tflops"], "peak_source": pk["src"] + " sustained bf16",
                           "traffic": BERT_GEMM_TRAFFIC_BYTES_PER_LAUNCH,
                           "traffic_source": "profiles/r01m_bert_ncu_full_summary.txt (bytes per launch, ncu --set full)",
                           "launches_per_step": n.value // prof_steps, "kernel_ms_per_step": tms.value / prof_steps,
                           "gemm_tflop_per_step": fl.value / prof_steps / 1e12}
    del net, model, optimizer
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------ VGG16 leg
VGG_BATCH, VGG_HW = 64, 224          # another_neural_net.py:64 batch size, :229/:240 CenterCrop(224)
VGG_GFLOP_PER_IMG_FWD = 30.95        # torchvision vgg16 at 224x224: 15.47 GMAC forward


def vgg_leg(dev, rank, world, steps, warmup, timed):
    """The reference's VGG16 recipe (another_neural_net.py:244-300): pretrained part frozen, classifier[6] replaced,
    optim.Adam(model.parameters()), train-mode forward (dropout in the 4096-wide layers) + backward through the head;
    then the eval-mode forward at batch 256.  Every rank runs its own replica on its own data (no collective)."""
    import torch
    import b200net
    from oracle import torch_oracle as O     # synthetic batch generator + head definition only
    torch.manual_seed(O.SEED_WEIGHTS)
    model = b200net.models.vgg16(pretrained=False)
    for param in model.parameters():
        param.requires_grad = False
    model.classifier[6] = O.make_vgg_head()
    model.to(dev).train()
    criterion = b200net.nn.NLLLoss()
    optimizer = b200net.optim.Adam(model.parameters())
    x, y = O.make_batch(VGG_BATCH, VGG_HW, 3000 + rank, dev)

    def step():
        optimizer.zero_grad()
        logps = model.forward(x)
        loss = criterion(logps, y)
        loss.backward()
        optimizer.step()
        return loss

    for _ in range(max(warmup, 3)):
        step()
    ms = timed(step, steps)
    model.eval()
    with torch.no_grad():
        xi = torch.rand(256, 3, VGG_HW, VGG_HW, device=dev)
        for _ in range(3):
            model(xi)
        ms_inf = timed(lambda: model(xi).argmax(1), 5)
    pk = peaks()
    train_ips = world * VGG_BATCH * steps / (ms * 1e-3)
    inf_ips = world * 256 * 5 / (ms_inf * 1e-3)
    del model, optimizer, xi
    torch.cuda.empty_cache()
    return {"workload": "vgg16 reference recipe (frozen pretrained part, Adam on classifier[6]), batch %d/GPU, 3x%dx%d; "
                        "eval forward at batch 256" % (VGG_BATCH, VGG_HW, VGG_HW),
            "train_img_s": train_ips, "train_ms_per_step": ms / steps, "infer_img_s": inf_ips,
            "infer_ms_per_batch": ms_inf / 5,
            "infer_tensor_util": VGG_GFLOP_PER_IMG_FWD * 1e9 * inf_ips / world / 1e12 / pk["tflops"]}


def run_bert_reference(args, rank):
    if rank != 0:
        return
    sps, ms, cores = bert_cpu_reference_run(args.steps, args.warmup, batch=8)
    sample = ("transformers BertForSequenceClassification + torch.optim.AdamW on stock torch %s CPU fp32 executing "
              "pytorch_on_language_distr.py:257-275; each step = 8 sequences of the %dx%d workload"
              % (__import__("torch").__version__, BERT_BATCH, BERT_SEQ))
    print(json.dumps({
        "impl": "reference", "metric": "bert_base_train_sequences_per_sec", "value": sps, "unit": "seq/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": BERT_WORKLOAD % (BERT_BATCH, BERT_SEQ), "cpu_sample_batch": 8},
        "cpu_baseline": {"value": sps, "unit": "seq/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": sps, "unit": "seq/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def hbm_kernels_leg(dev, model, optimizer):
    """Achieved HBM GB/s of the bandwidth kernels SURVEY 8(d) names, each timed alone with CUDA events on the largest
    activation of the step (layer1's block output at N=256: [409600 rows][256 channels] bf16 = 210 MB, > L2) and, for
    the optimizer, on the model's own 161 tensors.  Bytes are the ALGORITHMIC ones of SURVEY 8(d): apply 4 B/elem
    (+1/8 B mask), backward reduction 4 B/elem (+1/8), backward dx 6 B/elem (+1/8), SGD 12 B/param."""
    import torch
    from b200net import ops
    rows, C = 256 * 40 * 40, 256
    x = torch.randn(rows, C, device=dev).to(torch.bfloat16)
    dz = torch.randn(rows, C, device=dev).to(torch.bfloat16)
    gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    mean, invstd, scale, shift = ops.bn_train_stats(x, gamma, beta, None, None, 0.0, 1e-5)
    z, mask = ops.bn_apply(x, scale, shift, True, None, want_mask=True)
    elems = rows * C

    def t(fn, iters=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters * 1e-3

    out = {}
    pk = peaks()
    zbuf, mbuf = torch.empty_like(x), torch.empty_like(mask)
    lib = __import__("b200net")._C.lib
    st = torch.cuda.current_stream().cuda_stream
    sec = t(lambda: lib.b200_bn_apply(x.data_ptr(), None, zbuf.data_ptr(), mbuf.data_ptr(), rows, C, 1, scale.data_ptr(),
                                      shift.data_ptr(), 1, 0, st))
    out["bn_apply_relu_mask"] = {"gbs": elems * 4.125 / sec / 1e9, "bytes_per_elem": 4.125}
    sec_all = t(lambda: ops.bn_backward(dz, None, x, gamma, mean, invstd, True, want_dres=False, mask=mask))
    # the two passes of the backward share one entry point: time the reduction alone through the statistics kernel
    # of the same structure (two inputs -> partial sums) is not exposed, so report the pair against its 10.25 B/elem
    out["bn_backward_reduce_plus_dx"] = {"gbs": elems * 10.25 / sec_all / 1e9, "bytes_per_elem": 10.25}
    sec = t(lambda: ops.bn_train_stats(x, gamma, beta, None, None, 0.0, 1e-5))
    out["bn_stats"] = {"gbs": elems * 2.0 / sec / 1e9, "bytes_per_elem": 2.0}
    params = [p for p in model.parameters() if p.grad is not None]
    nparam = sum(p.numel() for p in params)
    sec = t(lambda: optimizer.step(), iters=5)
    out["mt_sgd"] = {"gbs": nparam * 12.0 / sec / 1e9, "bytes_per_param": 12.0, "params": nparam}
    for v in out.values():
        v["frac_of_hbm_peak"] = v["gbs"] / pk["hbm"]
    out["peak_gbs"] = pk["hbm"]
    out["tensor"] = "[%d rows][%d ch] bf16 (%.0f MB), CUDA events, 10 launches each" % (rows, C, elems * 2 / 1e6)
    del x, dz, z, mask, zbuf, mbuf
    torch.cuda.empty_cache()
    return out


def argmax_flips(net, dev, n=256):
    """bf16 eval predictions against the fp32 CPU oracle on n images of the inference batch shape: how many argmax
    values differ, and the oracle's largest top-2 margin among those rows (flips only happen at near-ties)."""
    import torch
    from oracle import torch_oracle as O
    ref = O.build_reference_model(dropout=False)
    ref.load_state_dict({k: v.detach().cpu() for k, v in net.state_dict().items()})
    ref.eval()
    torch.set_num_threads(O.cpu_threads())
    x, _ = O.make_batch(n, HW, 4242)
    with torch.no_grad():
        lo = ref(x).double()
        lg = net(x.to(dev)).double().cpu()
    srt = lo.sort(dim=1, descending=True).values
    margin = srt[:, 0] - srt[:, 1]
    fl = lo.argmax(1) != lg.argmax(1)
    return {"flips": int(fl.sum()), "of": n, "checker": "oracle (stock torch CPU fp32), same weights",
            "max_oracle_margin_among_flipped": float(margin[fl].max()) if fl.any() else 0.0,
            "median_oracle_margin": float(margin.median()), "max_abs_logprob_dev": float((lg - lo).abs().max())}


def ddp_check(model, criterion, x_dev, y_dev, world):
    """Driver-run correctness of the multi-GPU path (the GPU test box has one GPU): after the timed region,
    (a) every parameter is bit-identical on all ranks, (b) the bucket gradients of a synchronised backward equal the
    mean over ranks of the gradients of an unsynchronised (no_sync) backward of the same batch."""
    import torch
    import torch.distributed as dist
    net = model.module
    worst_param = 0.0
    for p in net.parameters():
        ref = p.detach().clone()
        dist.broadcast(ref, 0)
        worst_param = max(worst_param, float((ref - p.detach()).abs().max()))
    from b200net import functional as BF
    drop0 = BF._DROPOUT_CALLS[0]           # both backward passes must draw the same dropout mask in the head
    model.zero_grad(set_to_none=True)
    with model.no_sync():
        criterion(model(x_dev), y_dev).backward()
    local = [p.grad.detach().clone() for p in net.parameters()]
    for g in local:
        dist.all_reduce(g, op=dist.ReduceOp.SUM)
        g.div_(world)
    model.zero_grad(set_to_none=True)
    BF._DROPOUT_CALLS[0] = drop0
    criterion(model(x_dev), y_dev).backward()
    torch.cuda.synchronize()
    worst = 0.0
    for p, g in zip(net.parameters(), local):
        worst = max(worst, float(((p.grad - g).abs().max() / g.abs().max().clamp_min(1e-20))))
    model.zero_grad(set_to_none=True)
    t = torch.tensor([worst_param, worst], device=x_dev.device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return {"params_identical_across_ranks": bool(t[0].item() == 0.0), "max_param_diff": t[0].item(),
            "bucket_grad_vs_mean_of_local_max_rel": t[1].item(), "ok": bool(t[0].item() == 0.0 and t[1].item() < 1e-4)}


def comm_leg(model, dev, world, timed, step_sync, step_nosync, steps):
    """All-reduce bus bandwidth on the wrapper's own buckets (stand-alone, device-timed) and the part of a step that
    data parallelism adds: the DDP step minus the SAME step of an unwrapped replica in the same process (both replay
    CUDA graphs; every rank runs its replica at the same time, so host and memory contention are alike)."""
    import torch
    import torch.distributed as dist
    sizes = [b.flat.numel() * 4 for b in model.buckets]
    total = sum(sizes)
    def allreduce_all():
        for b in model.buckets:
            dist.all_reduce(b.flat, op=dist.ReduceOp.AVG)
    scratch = [b.flat.clone() for b in model.buckets]
    for _ in range(3):
        allreduce_all()
    ms = timed(allreduce_all, 10) / 10
    for b, s0 in zip(model.buckets, scratch):
        b.flat.copy_(s0)
    busbw = total * 2.0 * (world - 1) / world / (ms * 1e-3) / 1e9
    t_sync = timed(step_sync, steps) / steps
    t_nosync = timed(step_nosync, steps) / steps
    return {"buckets": len(sizes), "bytes": total, "allreduce_ms_all_buckets_standalone": ms, "busbw_gbs": busbw,
            "step_ms_ddp": t_sync, "step_ms_unwrapped_replica": t_nosync, "exposed_comm_ms": t_sync - t_nosync}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200net", choices=["b200net", "reference"])
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--batch", type=int, default=PER_GPU_BATCH)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="resnet", choices=["resnet", "bert"],
                    help="resnet = the headline line (BASELINE.json metric); bert = a full line for configs[4]")
    ap.add_argument("--no-bert", action="store_true", help="skip the BERT leg of the default line")
    ap.add_argument("--no-vgg", action="store_true", help="skip the VGG16 leg of the default line")
    ap.add_argument("--local_rank", type=int, default=0)     # torch.distributed.launch compatibility
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(args.local_rank)))

    if args.impl == "reference":
        return run_bert_reference(args, rank) if args.workload == "bert" else run_reference(args, rank)

    import torch
    import torch.distributed as dist
    import b200net
    from oracle import torch_oracle as O   # data generator + head definition only (checker infrastructure)

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (b200net has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    distributed = world > 1
    if distributed:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    b200net.set_precision(args.precision)

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if distributed:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item()

    if args.workload == "bert":
        if args.precision != "bf16":
            raise SystemExit("bench.py --workload bert: bf16 only")
        sampler = ClockSampler(local_rank)
        sampler.start()
        leg = bert_leg(dev, rank, world, distributed, args.steps, args.warmup, timed, with_roofline=True)
        clocks = sampler.stop()
        if rank == 0:
            out = {"metric": "bert_base_train_sequences_per_sec", "value": leg["seq_s"], "unit": "seq/s",
                   "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": leg["ms_per_step"],
                   "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
                   "data": "synthetic",
                   "config": {"workload": leg["workload"], "global_batch": world * BERT_BATCH, "seq_len": BERT_SEQ,
                              "parallelism": "dp%d" % world, "tokens_s": leg["tokens_s"],
                              "l2": "per-step working set (saved activations ~3 GB, 440 MB of weights) exceeds the "
                                    "126 MB L2; no explicit flush",
                              "tensor_util_step": leg["tensor_util_step"]},
                   "clocks": clocks, "e2e": leg["e2e"], "gpu_launches": leg["gpu_launches_per_step"] * args.steps,
                   "gpu_launches_per_step": leg["gpu_launches_per_step"], "roofline": leg["roofline"]}
            if world == 1 and not args.no_cpu_baseline:
                sps, cms, cores = bert_cpu_reference_run(steps=2, warmup=1, batch=8)
                out["cpu_baseline"] = {"value": sps, "unit": "seq/s", "cores": cores, "kind": "port",
                                       "sample": "oracle (transformers + stock torch CPU fp32, the reference's "
                                                 "train-step statements) on 8 sequences of length %d per step, "
                                                 "1 warm-up + 2 timed steps" % BERT_SEQ}
            print(json.dumps(out), flush=True)
        if distributed:
            dist.barrier()
            dist.destroy_process_group()
        return

    torch.manual_seed(O.SEED_WEIGHTS)
    model = b200net.models.resnet50(pretrained=False)
    model.fc = b200net.nn.Sequential(b200net.nn.Linear(2048, 512), b200net.nn.ReLU(), b200net.nn.Dropout(0.2),
                                     b200net.nn.Linear(512, 10), b200net.nn.LogSoftmax(dim=1))
    model.to(dev).train()
    criterion = b200net.nn.NLLLoss()
    optimizer = b200net.optim.SGD(model.parameters(), lr=0.001)
    if distributed:
        model = b200net.parallel.DistributedDataParallel(model, device_ids=[local_rank], output_device=local_rank)

    B = args.batch
    x_host, y_host = O.make_batch(B, HW, 1000 + rank)
    x_host, y_host = x_host.pin_memory(), y_host.pin_memory()
    x_dev, y_dev = x_host.to(dev), y_host.to(dev)

    def step_resident():
        optimizer.zero_grad()
        logps = model.forward(x_dev)
        loss = criterion(logps, y_dev)
        loss.backward()
        optimizer.step()
        return loss

    def step_e2e():
        # the reference loop body verbatim (...on_image.ipynb:667-673): H2D copy, step, loss.item()
        inputs, labels = x_host.to(dev, non_blocking=True), y_host.to(dev, non_blocking=True)
        optimizer.zero_grad()
        logps = model.forward(inputs)
        loss = criterion(logps, labels)
        loss.backward()
        optimizer.step()
        return loss.item()

    lib = b200net._C.lib
    # kernels per step, counted on one eager step (later steps replay the backbone as CUDA graphs: same kernels,
    # launched by the graph instead of one call each)
    step_resident()
    torch.cuda.synchronize()
    l0 = lib.b200_launch_count()
    step_resident()
    launches = lib.b200_launch_count() - l0
    for _ in range(max(args.warmup, 3) + 4):     # includes the eager calls that precede graph capture
        step_resident()
    sampler = ClockSampler(local_rank)
    sampler.start()
    total_ms = timed(step_resident, args.steps)
    clocks = sampler.stop()
    grecs = [r for r in b200net.functional._GRAPHS.values() if r != "failed"]
    graphs_used = {"backbone_graphs": len(grecs), "replays": int(sum(r.uses for r in grecs)),
                   "training_graph": any(r.bwd is not None and r.uses >= args.steps for r in grecs)}
    if args.precision == "bf16" and not graphs_used["training_graph"] and os.environ.get("B200_CUDA_GRAPHS", "1") != "0":
        raise SystemExit("bench.py: the timed region did not replay the training graph (capture failed?)")
    for _ in range(3):
        step_e2e()
    e2e_verbatim_ms = timed(step_e2e, args.steps)

    def e2e_loop():
        # the same loop with the loader wrapped by the library's prefetcher (b200net.utils.DevicePrefetcher): the copy
        # of batch i+1 overlaps step i; every step still moves its batch host->device and reads its loss back
        loader = b200net.utils.DevicePrefetcher(((x_host, y_host) for _ in range(args.steps)), dev)
        for inputs, labels in loader:
            inputs, labels = inputs.to(dev), labels.to(dev)
            optimizer.zero_grad()
            logps = model.forward(inputs)
            loss = criterion(logps, labels)
            loss.backward()
            optimizer.step()
            loss.item()

    # two loops of K steps, the faster one is reported: a single host hiccup (allocator, page cache) inside one 20-step
    # loop once halved this number on an otherwise idle box while `verbatim_loop_value` next to it was unaffected
    e2e_ms = min(timed(e2e_loop, 1), timed(e2e_loop, 1))

    # roofline leg: per-launch CUDA-event timing of the tcgen05 GEMM kernel over 3 extra steps (outside the timed region)
    import ctypes
    roof = None
    if args.precision == "bf16":
        prof_steps = 3
        torch.cuda.synchronize()
        b200net.set_cuda_graphs(False)            # per-launch events need the eager path
        lib.b200_gemm_timing_enable(1)
        for _ in range(prof_steps):
            step_resident()
        torch.cuda.synchronize()
        ms, n, fl = ctypes.c_double(), ctypes.c_int64(), ctypes.c_double()
        lib.b200_gemm_timing_read(ctypes.byref(ms), ctypes.byref(n), ctypes.byref(fl))
        lib.b200_gemm_timing_enable(0)
        b200net.set_cuda_graphs(True)
        pk = peaks()
        algo_flops = GFLOP_PER_IMG_TRAIN * 1e9 * B * prof_steps
        achieved = algo_flops / (ms.value * 1e-3) / 1e12
        roof = {"bound": "tensor", "kernel": "conv_sm100_kernel (tcgen05 implicit-GEMM conv fprop/dgrad/wgrad family)",
                "achieved": achieved, "peak": pk["tflops"], "unit": "TFLOP/s", "frac": achieved / pk["tflops"],
                "peak_source": pk["src"] + " sustained bf16", "traffic": _named_launch_traffic(),
                "algorithmic_bytes": CONV_ALGO_BYTES_PER_LAUNCH, "traffic_launch": CONV_TRAFFIC_LAUNCH,
                "traffic_source": "profiles/r02_ncu_named_launches.txt (ncu --set full, dram__bytes_read.sum + "
                                  "dram__bytes_write.sum of that one launch)",
                "per_layer": {"roofline_ms": conv_per_layer_roofline_ms(B, pk["tflops"], pk["hbm"]),
                              "measured_ms": ms.value / prof_steps,
                              "frac": conv_per_layer_roofline_ms(B, pk["tflops"], pk["hbm"]) / (ms.value / prof_steps),
                              "note": "sum over the step's conv launches of max(tensor time, HBM time) per launch "
                                      "(sustained bf16 peak, measured HBM copy rate) / measured kernel time of the "
                                      "same launches; `frac` above is the whole family against the tensor peak only"},
                "launches_per_step": n.value // prof_steps, "kernel_ms_per_step": ms.value / prof_steps,
                "executed_gemm_tflop_per_step": fl.value / prof_steps / 1e12}

    # inference leg (BASELINE.json configs[3] and the reference's per-image loop, ...on_image.ipynb:855-861):
    # eval mode, no_grad; every rank runs an independent replica on its own shard (no collective)
    infer = None
    if args.precision == "bf16":
        net = model.module if distributed else model
        net.eval()
        with torch.no_grad():
            xi = torch.rand(1024, 3, HW, HW, device=dev)
            for _ in range(6):
                net(xi)
            t_inf = timed(lambda: net(xi).argmax(1), 10)
            x1 = torch.rand(1, 3, 224, 224, device=dev)
            for _ in range(8):
                net(x1)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(100):
                int(net(x1).data.cpu().numpy().argmax())       # predict_image(): one D2H sync per image
            lat = (time.perf_counter() - t0) / 100 * 1e3
            flips = argmax_flips(net, dev) if rank == 0 else None
        net.train()
        infer = {"argmax_flips": flips, "batch": 1024, "img_s": world * 1024 * 10 / (t_inf * 1e-3), "ms_per_batch": t_inf / 10,
                 "batch1_224px_latency_ms": lat, "tensor_util": 4.1727e9 * 1024 * 10 / (t_inf * 1e-3) / 1e12 / peaks()["tflops"]}
        del xi

    hbm = None
    if args.precision == "bf16" and rank == 0 and world == 1:
        hbm = hbm_kernels_leg(dev, model, optimizer)
    ddpc = comm = None
    if distributed:
        ddpc = ddp_check(model, criterion, x_dev, y_dev, world)

        torch.manual_seed(O.SEED_WEIGHTS)
        plain = b200net.models.resnet50(pretrained=False)
        plain.fc = b200net.nn.Sequential(b200net.nn.Linear(2048, 512), b200net.nn.ReLU(), b200net.nn.Dropout(0.2),
                                         b200net.nn.Linear(512, 10), b200net.nn.LogSoftmax(dim=1))
        plain.to(dev).train()
        popt = b200net.optim.SGD(plain.parameters(), lr=0.001)

        def step_plain():
            b200net.functional.set_grad_sink(None)       # the wrapper installs its sink on every forward
            popt.zero_grad()
            loss = criterion(plain.forward(x_dev), y_dev)
            loss.backward()
            popt.step()
            return loss
        for _ in range(8):
            step_plain()
        for _ in range(3):
            step_resident()
        comm = comm_leg(model, dev, world, timed, step_resident, step_plain, 10)
        del plain, popt

    # BERT / IMDB leg (configs[4]): the same reference statements on the second model of the harness
    bert_out = None
    if args.precision == "bf16" and not args.no_bert:
        del x_dev
        torch.cuda.empty_cache()
        bert_out = bert_leg(dev, rank, world, distributed, 10, 3, timed, with_roofline=False)
    vgg_out = None
    if args.precision == "bf16" and not args.no_vgg:
        vgg_out = vgg_leg(dev, rank, world, 10, 3, timed)

    if rank == 0:
        ips = world * B * args.steps / (total_ms * 1e-3)
        e2e_ips = world * B * args.steps / (e2e_ms * 1e-3)
        out = {
            "metric": METRIC, "value": ips, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD % (B, HW, HW),
                       "global_batch": world * B, "parallelism": "dp%d" % world,
                       "l2": "per-step working set (saved activations ~6 GB) far exceeds the 126 MB L2; no explicit flush",
                       "cuda_graphs": "backbone forward/backward replayed as CUDA graphs after %d eager calls%s"
                                      % (b200net.functional.GRAPH_WARMUP,
                                         " (under DDP the bucket packs are inside the backward graph and the "
                                         "all-reduces follow the replay)" if distributed else ""),
                       "step_tflop": GFLOP_PER_IMG_TRAIN * B / 1e3,
                       "tensor_util_step": GFLOP_PER_IMG_TRAIN * 1e9 * ips / world / 1e12 / peaks()["tflops"]},
            "clocks": clocks,
            "e2e": {"value": e2e_ips, "unit": UNIT, "h2d_bytes_per_step": x_host.numel() * 4 + y_host.numel() * 8,
                    "d2h_bytes_per_step": 4, "ms_per_step": e2e_ms / args.steps,
                    "loader": "b200net.utils.DevicePrefetcher around the host batches (H2D of batch i+1 overlaps step i); best of two K-step loops",
                    "verbatim_loop_value": world * B * args.steps / (e2e_verbatim_ms * 1e-3)},
            "gpu_launches": int(launches) * args.steps,
            "gpu_launches_per_step": int(launches),
            "graphs_used": graphs_used,
        }
        if hbm:
            out["hbm_kernels"] = hbm
        if ddpc:
            out["ddp_check"] = ddpc
        if comm:
            out["comm"] = comm
        if roof:
            out["roofline"] = roof
        if infer:
            out["infer"] = infer
        if bert_out:
            out["bert"] = bert_out
        if vgg_out:
            out["vgg16"] = vgg_out
        if world == 1 and not args.no_cpu_baseline:
            cips, cms, cores = cpu_reference_run(steps=5, warmup=2)
            out["cpu_baseline"] = {"value": cips, "unit": UNIT, "cores": cores, "kind": "port",
                                   "sample": "oracle (stock torch CPU fp32, the reference's train-step statements) on "
                                             "%d images per step of the same workload, 2 warm-up + 5 timed steps"
                                             % CPU_SAMPLE_BATCH}
        print(json.dumps(out), flush=True)
    if distributed:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
