"""BASELINE.json configs[4] / SURVEY.md section 8(d) config 5 and 8(f) row 3: the generator feeding the reference's
trainer inline.

The sink is the reference's model by shape -- ``PeraNet`` (models.py:21-146, 210-253): stock torchvision ResNet-18 with
``fc`` replaced by Identity, forward hooks on layer2 / layer3, global average pools concatenated (128 + 256 + 512),
``concatenator`` Linear + BN, a 4-layer ``latent_space`` MLP and a Linear classifier; cross-entropy, SGD(momentum 0.9,
weight decay 5e-4) as ``configure_optimizers`` (models.py:336-341).  The model is stock PyTorch / cuDNN by contract
(out of scope for hand-written kernels); random initialisation (no checkpoints on the box).
What this module measures is the hand-off: ``ssad_generate`` writes the normalised bf16 NCHW batch and the labels on
the SAME stream the training step runs on -- no host round trip, no DataLoader -- and how much of a training step the
generation costs.  One process per GPU; with more than one rank the model is wrapped in DistributedDataParallel
(the only collective: the gradient all-reduce of the 12 M parameters over NCCL / NVLink).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


class PeraNetShaped(nn.Module):
    """The reference's ``PeraNet`` forward (models.py:210-253, image-level branch) without Lightning."""

    def __init__(self, latent_layers: int = 5, base_dim: int = 512, num_classes: int = 4):
        super().__init__()
        from torchvision import models
        self.feature_extractor = models.resnet18(weights=None)           # random init: no checkpoint files on the box
        self.feature_extractor.fc = nn.Identity()                         # models.py:59-63
        self.activations = {}
        self.feature_extractor.layer2.register_forward_hook(lambda m, i, o: self.activations.__setitem__("layer2", o))
        self.feature_extractor.layer3.register_forward_hook(lambda m, i, o: self.activations.__setitem__("layer3", o))
        dim = 512 + 128 + 256                                             # models.py:124-136
        self.concatenator = nn.Sequential(nn.Linear(dim, base_dim, bias=False), nn.BatchNorm1d(base_dim))
        layers = []
        for _ in range(latent_layers - 1 - 1):                            # models.py:66-88 with latent_space_layers - 1
            layers.append(nn.Sequential(nn.Linear(base_dim, base_dim, bias=False), nn.BatchNorm1d(base_dim), nn.ReLU(inplace=True)))
        layers += [nn.Linear(base_dim, 512, bias=True), nn.BatchNorm1d(512)]
        self.latent_space = nn.Sequential(*layers)
        self.classifier = nn.Linear(512, num_classes)

    def forward(self, x):
        self.activations = {}
        feats = torch.flatten(self.feature_extractor(x), 1)
        f2 = torch.flatten(F.adaptive_avg_pool2d(self.activations["layer2"], (1, 1)), 1)
        f3 = torch.flatten(F.adaptive_avg_pool2d(self.activations["layer3"], (1, 1)), 1)
        feats = torch.cat([f2, torch.cat([f3, feats], dim=1)], dim=1)
        emb = self.latent_space(self.concatenator(feats))
        return self.classifier(emb), emb


def _timed(fn, steps, warmup, rt):
    dev, world, dist = rt["dev"], rt["world"], rt["dist"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
    for i in range(warmup):
        fn(i)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms / steps


def measure(args, rt, make_workload, steps: int = 20, warmup: int = 5, batch: int = 512) -> dict:
    """Per-GPU batch of 512 x 256x256 image-level items -> bf16 NCHW + labels -> one SGD step of the ResNet-18-shaped
    model, all on one stream.  Returns step times with / without the generation and the generator's share."""
    from .generator import BatchGenerator
    from .params import sample_params
    dev, world, rank = rt["dev"], rt["world"], rt["rank"]
    cfg, images, mask, donors, params, name, ab = make_workload(5, batch, seed=rank)
    gen = BatchGenerator(cfg, dev, mask=mask, donors=donors)
    src = torch.from_numpy(images).to(dev)
    nsets = 4                                # parameter sets drawn ahead (fresh records per step, cycling)
    rng = np.random.default_rng(4321 + rank)
    psets = [params] + [sample_params(cfg, batch, rng) for _ in range(nsets - 1)]
    x = torch.empty((batch, 3, cfg.work_size[1], cfg.work_size[0]), dtype=torch.bfloat16, device=dev)
    y = torch.empty((batch,), dtype=torch.int64, device=dev)
    plans = [gen.plan(src, p, want_u8=False, norm_dtype=torch.bfloat16, out_norm=x, y_out=y) for p in psets]
    if getattr(args, "graph", 1):
        for pl in plans:
            gen.capture(pl)

    torch.manual_seed(1234)
    model = PeraNetShaped().to(dev).to(memory_format=torch.channels_last)
    net = model
    if world > 1:
        from torch.nn.parallel import DistributedDataParallel
        net = DistributedDataParallel(model, device_ids=[dev.index])
    opt = torch.optim.SGD(net.parameters(), 0.03, momentum=0.9, weight_decay=0.0005)     # models.py:336-337

    def train_step(_i):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            y_hat, _ = net(x.contiguous(memory_format=torch.channels_last))
            loss = F.cross_entropy(y_hat.float(), y)                       # models.py:260
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)

    def gen_step(i):
        gen.run(plans[i % nsets])                                          # writes x, y on the current stream

    def both(i):
        gen_step(i)
        train_step(i)

    gen_step(0)
    t_gen = _timed(gen_step, steps, warmup, rt)
    t_train = _timed(train_step, steps, warmup, rt)
    t_both = _timed(both, steps, warmup, rt)
    launches = int(gen.lib.ssad_generate_launches(plans[0].args[0], int(bool(getattr(args, "graph", 1)))))
    return {"workload": name, "batch_per_gpu": batch, "global_batch": batch * world, "steps": steps,
            "value": world * batch / (t_both * 1e-3), "unit": "images/s (generation + ResNet-18 training step, inline)",
            "ms_per_step_with_generation": t_both, "ms_per_step_train_only": t_train, "generate_ms": t_gen,
            "generator_share_of_step": t_gen / t_both, "generator_images_per_s_per_gpu": batch / (t_gen * 1e-3),
            "roofline_frac_generator": (batch * ab / (t_gen * 1e-3) / 1e9),
            "gpu_launches_generator": steps * launches,
            "model": "ResNet-18 + layer2/3 hooks + concatenator + latent MLP + classifier (models.py:21-146, 210-253), random init, "
                     "bf16 autocast, channels_last, SGD momentum 0.9" + (", DistributedDataParallel (NCCL)" if world > 1 else ""),
            "precision": "bf16 autocast (the reference trains with precision=16, tools.py:260-268)"}


def bench_line(args, rt, make_workload, config_block, clock_sampler) -> dict:
    """The JSON line of ``bench.py --config 5``."""
    import json
    import os
    clock_sampler.start()
    steps = args.steps or 30
    m = measure(args, rt, make_workload, steps=steps, warmup=max(args.warmup, 3), batch=args.batch or 512)
    clocks = clock_sampler.stop()
    peak = 6650.0
    try:
        peak = float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    world = rt["world"]
    ach = m.pop("roofline_frac_generator")
    return {"metric": "synthetic_anomaly_images_per_sec", "value": m["value"], "unit": "images/s", "n_gpus": world, "steps": steps,
            "warmup": max(args.warmup, 3), "ms_per_step": m["ms_per_step_with_generation"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": config_block(5, m["workload"], m["batch_per_gpu"], world, "bf16", 1, bool(args.graph)),
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                         "kernel": "ssad_generate alone (the training step is stock cuDNN, outside the roofline claim)"},
            "e2e": None, "gpu_launches": m["gpu_launches_generator"], "trainer_feed": m, "clocks": clocks}
