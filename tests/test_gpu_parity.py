"""GPU parity (run on the B200 box): the CUDA path through the C ABI against the CPU oracle on the
same seeded inputs and parameters, and against the golden vectors made from the reference's own code.

Bars: labels, mask decisions, paste positions and every pasted / warped / rotated pixel bit-exact;
jittered pixels bit-exact as well (the contract allows +-1 LSB; the kernels reproduce Pillow's
float32 sequence, so any deviation is reported); normalised f32 output bit-exact against
torchvision's ToTensor+Normalize of the oracle image; bf16 = round-to-nearest-even of that."""
import os

import numpy as np
import pytest
import torch

import scenarios as S
from golden import make_golden as G

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _reload_env():
    from ssad_b200 import _native
    _native.load().ssad_reload_env()


@pytest.fixture(autouse=True)
def _default_env_after_each_test():
    yield
    for k in [k for k in os.environ if k.startswith("SSAD_V7_")]:
        del os.environ[k]
    _reload_env()


def _check_against_oracle(name, batch, monkeypatch=None, env=None):
    import gpu_utils as U
    from oracle import generator as og
    from PIL import Image
    sc = S.build(name, batch)
    if env:
        for k, v in env.items():
            monkeypatch.setenv(k, v)
    _reload_env()                                # the library reads its tuning environment once
    out = U.run_cuda(sc)
    x = out["x_u8"].cpu().numpy()
    y = out["y"].cpu().numpy()
    xn = out["x"].cpu().numpy()
    dbg = out["debug"]
    for i in range(len(sc.params)):
        ox, oy, info = S.run_oracle(sc, i)
        assert int(y[i]) == oy, (name, i)
        if oy > 0:
            assert int(dbg["n_coords"][i]) == info["n_coords"]
        if oy in (1, 2):
            assert bool(dbg["sim"][i]) == info["sim"], (name, i, dbg["cosine"][i], info["cos"])
            assert abs(dbg["cosine"][i] - info["cos"]) < 1e-12
            got = [tuple(int(v) for v in dbg["paste"][i][k]) for k in range(len(info["paste"]))]
            assert got == [tuple(int(v) for v in p) for p in info["paste"]], (name, i)
        diff = np.abs(x[i].astype(np.int16) - ox.astype(np.int16))
        assert diff.max() == 0, (name, i, int(diff.max()), int((diff > 0).sum()))
        ref_n = og.to_tensor_normalize(Image.fromarray(ox)).numpy()
        assert np.array_equal(xn[i], ref_n), (name, i)


@pytest.mark.parametrize("name", S.ALL)
def test_generate_matches_oracle(name):
    _check_against_oracle(name, 12)


@pytest.mark.parametrize("name", S.BIG)
def test_generate_big_images_match_oracle(name):
    """512 / 768 / 1024-pixel images: several slabs per image, taller canvases (two canvas CTAs per image at 1024)."""
    _check_against_oracle(name, 5)


@pytest.mark.parametrize("env", [{"SSAD_V7_NOFAST": "1"}, {"SSAD_V7_UPT": "3"}, {"SSAD_V7_NC": "1"}, {"SSAD_V7_NC": "16"}, {"SSAD_V7_SPLIT": "3"},
                                 {"SSAD_V7_SPLIT": "8", "SSAD_V7_NOPDL": "1"}, {"SSAD_V7_SPLIT": "4", "SSAD_V7_LANES": "1"},
                                 {"SSAD_V7_SPLIT": "7", "SSAD_V7_LANES": "3"}])
@pytest.mark.parametrize("name", ["image_level", "image_level_256", "patch_level", "non_fixed", "texture"])
def test_v7_loop_and_grid_variants_match_oracle(name, env, monkeypatch):
    """v7's general unit loops (the fast ones need a power-of-two width and whole slabs), odd slab sizes, and the
    number of CTAs that share an image's patch canvas, and the batch cut into chunks that run on lanes (helper streams;
    with more chunks than lanes a lane's workspace is reused by its next chunk)."""
    _check_against_oracle(name, 12, monkeypatch, env)


@pytest.mark.parametrize("name,batch", [("patch_level", 257), ("texture", 259), ("image_level_256", 3)])
def test_default_chunking_matches_oracle(name, batch):
    """Batches past the chunking threshold (256) with the library's own choice of lanes and chunks (odd sizes: the last
    chunk is shorter), and a tiny batch, against the oracle image by image."""
    _check_against_oracle(name, batch)


@pytest.mark.parametrize("name", S.ALL)
def test_generate_matches_reference_golden(name):
    import gpu_utils as U
    g = np.load(os.path.join(GOLD, f"{name}.npz"))
    sc = S.build(name, G.BATCH[name])
    assert bytes(g["sha"]).hex() == G.scenario_sha(sc)
    out = U.run_cuda(sc)
    assert np.array_equal(out["y"].cpu().numpy(), g["y"])
    assert np.array_equal(out["x_u8"].cpu().numpy(), g["x"])
    assert np.array_equal(out["x"].cpu().numpy()[:2], g["xn"])


def test_bf16_output_is_rounded_f32():
    import gpu_utils as U
    sc = S.build("image_level", 6)
    a = U.run_cuda(sc, norm_dtype=torch.float32)["x"]
    b = U.run_cuda(sc, norm_dtype=torch.bfloat16)["x"]
    assert torch.equal(a.to(torch.bfloat16), b)


def test_no_jitter_no_warp_label0_is_identity():
    import gpu_utils as U
    sc = S.build("non_fixed", 4)
    sc.cfg.jitter = False
    sc.params = S.sample_params(sc.cfg, 4, np.random.default_rng(3), labels=np.zeros(4, dtype=np.int64))
    out = U.run_cuda(sc, norm_dtype=None)
    assert np.array_equal(out["x_u8"].cpu().numpy(), sc.images)


def test_large_batch_properties_config2_shape():
    """BASELINE config 2 at full size (512 x 256x256 windows): size-independent properties --
    label-0 images equal jitter(window) exactly; defect images differ from jitter-only output
    only inside the paste boxes; y follows the coverage rule."""
    import gpu_utils as U
    from ssad_b200 import synth
    B, S_, ps = 512, 288, 256
    cfg = S.GeneratorConfig(subject="bottle", imsize=(S_, S_), patch_localization=True, patch_size=ps,
                            patch_area_ratio=(0.1, 0.2), scar_area_ratio=(0.01, 0.02))
    base = synth.object_images(8, S_, S_, 5)
    imgs = base[np.arange(B) % 8]
    mask = synth.disc_mask(S_, S_, 0.5)
    params = S.sample_params(cfg, B, np.random.default_rng(9))
    sc = S.Scenario("cfg2", cfg, imgs, mask, None, params)
    gen = U.make_generator(sc)
    out = U.run_cuda(sc, norm_dtype=None, gen=gen)
    # same parameters with every label forced to 0 -> jitter only
    draws0 = {k: v.copy() for k, v in params.draws.items()}
    draws0["label"][:] = 0
    sc0 = S.Scenario("cfg2_0", cfg, imgs, mask, None, S.ssad_b200.params_from_draws(cfg, draws0))
    out0 = U.run_cuda(sc0, norm_dtype=None, gen=gen)
    x, x0 = out["x_u8"].cpu().numpy(), out0["x_u8"].cpu().numpy()
    y = out["y"].cpu().numpy()
    dbg = out["debug"]
    rec = params.records
    n_def = 0
    for i in range(B):
        cov_ok = 3 * int(S.window_mask(sc, i).sum()) >= cfg.coverage_min3
        assert y[i] == (params.draws["label"][i] if cov_ok else 0)
        changed = np.argwhere((x[i] != x0[i]).any(axis=2))
        if y[i] == 0:
            assert len(changed) == 0
            continue
        n_def += 1
        if len(changed) == 0:
            continue
        # jitter is pointwise except contrast (global mean): compare supports only when contrast mean is unchanged
        if dbg["contrast_mean"][i] != out0["debug"]["contrast_mean"][i]:
            continue
        k = 1 if y[i] == 1 else int(rec["scar_k"][i])
        w, h = (rec["rect"][i][2], rec["rect"][i][3]) if y[i] == 1 else rec["scar_size"][i]
        inside = np.zeros(len(changed), dtype=bool)
        for j in range(k):
            l, t = dbg["paste"][i][j]
            inside |= (changed[:, 1] >= l) & (changed[:, 1] < l + w) & (changed[:, 0] >= t) & (changed[:, 0] < t + h)
        assert inside.all()
    assert n_def > B // 3


@pytest.mark.parametrize("split", ["1", "3"])
@pytest.mark.parametrize("name", ["image_level_256", "patch_level", "texture"])
def test_graph_replay_matches_plain_launch(name, split, monkeypatch):
    """BatchGenerator.capture(): the whole ssad_generate call (kernels on the caller's stream and on the helper
    streams of the chunked pipeline) recorded in a CUDA graph; replays on fresh inputs equal plain launches."""
    import gpu_utils as U
    monkeypatch.setenv("SSAD_V7_SPLIT", split)
    _reload_env()
    sc = S.build(name, 12)
    gen = U.make_generator(sc)
    src = torch.from_numpy(sc.images).cuda()
    ref = gen.generate(src, sc.params, masks=U.per_image_masks(sc), want_u8=True, norm_dtype=torch.float32)
    torch.cuda.synchronize()
    buf = torch.zeros_like(src)
    plan = gen.plan(buf, sc.params, masks=U.per_image_masks(sc), want_u8=True, norm_dtype=torch.float32)
    gen.capture(plan)
    for _ in range(2):
        plan.out_u8.zero_(); plan.out_norm.zero_(); plan.y_out.zero_()
        buf.copy_(src)
        gen.run(plan)
        torch.cuda.synchronize()
        assert torch.equal(plan.out_u8, ref["x_u8"])
        assert torch.equal(plan.out_norm, ref["x"])
        assert torch.equal(plan.y_out, ref["y"])


@pytest.mark.parametrize("chunks", [1, 3])
@pytest.mark.parametrize("name", ["image_level_256", "patch_level", "non_fixed"])
def test_host_buffer_pipeline_matches_device_call(name, chunks):
    """generate_from_host: pinned host images in, host results out, the batch pipelined through copy / generate /
    copy streams -- same bytes as the device-resident call."""
    import gpu_utils as U
    from ssad_b200.generator import pack_mask_bits
    sc = S.build(name, 12)
    gen = U.make_generator(sc)
    ref = U.run_cuda(sc, norm_dtype=None, debug=False, gen=gen)
    h_src = torch.from_numpy(sc.images).pin_memory()
    W, H = sc.cfg.work_size
    h_out = torch.zeros((12, H, W, 3), dtype=torch.uint8).pin_memory()
    h_y = torch.zeros((12,), dtype=torch.int64).pin_memory()
    masks_fn = None
    if sc.mask is not None and sc.mask.ndim == 3:
        all_masks = U.per_image_masks(sc)
        state = {"lo": 0}

        def masks_fn(d_src):                     # the scenario's own masks, chunk by chunk
            lo = state["lo"]; state["lo"] = lo + d_src.shape[0]
            return all_masks[lo:lo + d_src.shape[0]]
    gen.generate_from_host(h_src, sc.params, h_out, h_y, chunks=chunks, masks_fn=masks_fn)
    assert torch.equal(h_out, ref["x_u8"].cpu())
    assert torch.equal(h_y, ref["y"].cpu())


@pytest.mark.parametrize("name", ["patch_level", "patch_level_32", "lines_patch"])
def test_host_buffer_pipeline_window_upload_matches_full_upload(name):
    """generate_from_host(upload="windows"): only the working window and the cut rectangle of every image cross PCIe
    (ssad_upload_windows reads the pinned host images directly).  Same bytes as the device-resident call -- also when
    the device staging buffer holds OTHER images from an earlier call everywhere else (a read outside the uploaded
    regions would show)."""
    import gpu_utils as U
    sc = S.build(name, 24)
    other = S.build(name, 24, seed=5)
    gen = U.make_generator(sc)
    ref = U.run_cuda(sc, norm_dtype=None, debug=False, gen=gen)
    W, H = sc.cfg.work_size
    h_out = torch.zeros((24, H, W, 3), dtype=torch.uint8).pin_memory()
    h_y = torch.zeros((24,), dtype=torch.int64).pin_memory()
    # an earlier call leaves the other scenario's images in the staging buffer
    gen.generate_from_host(torch.from_numpy(255 - other.images).pin_memory(), other.params, h_out, h_y, chunks=3, upload="full")
    h_src = torch.from_numpy(sc.images).pin_memory()
    for mode in ("windows", "full"):
        h_out.zero_(); h_y.zero_()
        gen.generate_from_host(h_src, sc.params, h_out, h_y, chunks=3, upload=mode)
        assert torch.equal(h_out, ref["x_u8"].cpu()), mode
        assert torch.equal(h_y, ref["y"].cpu()), mode


def test_window_upload_is_refused_where_the_whole_image_is_needed():
    import gpu_utils as U
    sc = S.build("image_level_256", 12)          # pre-warp: gathers from the whole image
    gen = U.make_generator(sc)
    W, H = sc.cfg.work_size
    h_out = torch.zeros((12, H, W, 3), dtype=torch.uint8).pin_memory()
    h_y = torch.zeros((12,), dtype=torch.int64).pin_memory()
    with pytest.raises(ValueError, match="windows"):
        gen.generate_from_host(torch.from_numpy(sc.images).pin_memory(), sc.params, h_out, h_y, upload="windows")


@pytest.mark.gpu
def test_host_buffer_pipeline_two_batches_in_flight():
    """generate_from_host(wait=False): a second call is issued while the first batch is still on its way; both deliver
    the bytes of the device-resident call once their events complete (the staging buffers are shared between calls)."""
    import gpu_utils as U
    scs = [S.build("patch_level", 24, seed=0), S.build("patch_level", 24, seed=7)]
    gen = U.make_generator(scs[0])               # one generator (one shared mask) serves both batches
    W, H = scs[0].cfg.work_size
    refs, outs, events = [], [], []
    for sc in scs:
        r = U.run_cuda(sc, norm_dtype=None, debug=False, gen=gen)
        refs.append({"x_u8": r["x_u8"].cpu().clone(), "y": r["y"].cpu().clone()})
    for k, sc in enumerate(scs):
        h_src = torch.from_numpy(sc.images).pin_memory()
        h_out = torch.zeros((24, H, W, 3), dtype=torch.uint8).pin_memory()
        h_y = torch.zeros((24,), dtype=torch.int64).pin_memory()
        events.append(gen.generate_from_host(h_src, sc.params, h_out, h_y, chunks=3, wait=False))
        outs.append((h_src, h_out, h_y))
    for k in range(2):
        events[k].synchronize()
        assert torch.equal(outs[k][1], refs[k]["x_u8"])
        assert torch.equal(outs[k][2], refs[k]["y"])
    assert not torch.equal(outs[0][1], outs[1][1])


@pytest.mark.parametrize("name", ["image_level_256", "patch_level", "texture", "non_fixed"])
def test_indexed_sources_match_gathered_batch(name):
    """SsadBatchDesc.src_indexed: the records carry the index of their source image in a resident set (no gather of the
    batch, datasets.py:211 becomes an index; per-image masks are then per source image) -- same bytes as running the
    gathered batch."""
    import gpu_utils as U
    sc = S.build(name, 12)
    gen = U.make_generator(sc)
    ref = U.run_cuda(sc, norm_dtype=torch.float32, debug=False, gen=gen)
    rng = np.random.default_rng(4)
    perm = rng.permutation(12)
    resident = torch.from_numpy(sc.images[perm]).cuda()           # resident[k] = image perm[k]; item i reads row inv[i]
    inv = np.argsort(perm)
    masks = U.per_image_masks(sc)
    if masks is not None:
        masks = masks[torch.as_tensor(perm).cuda()].contiguous()
    out = gen.generate(resident, sc.params, masks=masks, want_u8=True, norm_dtype=torch.float32, src_index=inv)
    torch.cuda.synchronize()
    assert torch.equal(out["x_u8"], ref["x_u8"])
    assert torch.equal(out["x"], ref["x"])
    assert torch.equal(out["y"], ref["y"])
    # a batch larger than the resident set (duplicated filenames, functional.py:64-68) and the chunked path
    big = np.concatenate([inv] * 25)[:290]
    pb = S.ssad_b200.params_from_draws(sc.cfg, {k: np.concatenate([v] * 25)[:290] for k, v in sc.params.draws.items()})
    out2 = gen.generate(resident, pb, masks=masks, want_u8=True, src_index=big)
    torch.cuda.synchronize()
    assert torch.equal(out2["x_u8"][:12], ref["x_u8"]) and torch.equal(out2["x_u8"][276:288], ref["x_u8"])


@pytest.mark.parametrize("chunks", [1, 3])
@pytest.mark.parametrize("name", ["image_level_256", "patch_level", "texture", "non_fixed", "lines_image"])
def test_resident_sources_to_host_match_device_call(name, chunks):
    """generate_resident_to_host: the source set stays on the device, per call only the records go up and the results
    come down (the DataModule's deployment) -- same bytes as the device call on the gathered batch, twice in a row."""
    import gpu_utils as U
    sc = S.build(name, 24)
    gen = U.make_generator(sc)
    ref = U.run_cuda(sc, norm_dtype=None, debug=False, gen=gen)
    perm = np.random.default_rng(4).permutation(24)
    resident = torch.from_numpy(sc.images[perm]).cuda()
    inv = np.argsort(perm)
    masks = U.per_image_masks(sc)
    if masks is not None:
        masks = masks[torch.as_tensor(perm).cuda()].contiguous()
    W, H = sc.cfg.work_size
    h_out = torch.zeros((24, H, W, 3), dtype=torch.uint8).pin_memory()
    h_y = torch.zeros((24,), dtype=torch.int64).pin_memory()
    for _ in range(2):
        h_out.zero_(); h_y.zero_()
        gen.generate_resident_to_host(resident, inv, sc.params, h_out, h_y, masks=masks, chunks=chunks)
        assert torch.equal(h_out, ref["x_u8"].cpu())
        assert torch.equal(h_y, ref["y"].cpu())


def test_two_host_threads_with_their_own_contexts():
    """ssad_context_create: two host threads drive the same device at once, each through its own context (its own helper
    streams); both get the bytes of a single-threaded call."""
    import threading
    import gpu_utils as U
    scs = [S.build("patch_level", 300, seed=1), S.build("image_level_256", 280, seed=2)]
    gens = [U.make_generator(sc) for sc in scs]
    refs = [U.run_cuda(sc, norm_dtype=None, debug=False, gen=g)["x_u8"].clone() for sc, g in zip(scs, gens)]
    outs, errs = [None, None], []

    def work(k):
        try:
            torch.cuda.set_device(0)
            ctx = gens[k].new_context()
            stream = torch.cuda.Stream()
            src = torch.from_numpy(scs[k].images).cuda()
            with torch.cuda.stream(stream):
                for _ in range(6):
                    o = gens[k].generate(src, scs[k].params, masks=U.per_image_masks(scs[k]), want_u8=True, context=ctx)
                stream.synchronize()
            outs[k] = o["x_u8"]
            gens[k].free_context(ctx)
        except Exception as e:           # noqa: BLE001
            errs.append(e)
    ts = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs
    for k in range(2):
        assert torch.equal(outs[k], refs[k])


def test_saturation_factors_up_to_two_match_oracle():
    """The table after Saturation serves factors up to 2 (torchvision ColorJitter(saturation <= 1)); beyond that the
    library refuses the batch instead of falling back to another kernel."""
    import gpu_utils as U
    sc = S.build("image_level", 12)
    draws = {k: v.copy() for k, v in sc.params.draws.items()}
    draws["jit_s"] = np.linspace(0.0, 2.0, 12).astype(np.float32).astype(np.float64)
    draws["jit_b"] = np.linspace(0.5, 1.5, 12).astype(np.float32).astype(np.float64)
    sc.params = S.ssad_b200.params_from_draws(sc.cfg, draws)
    out = U.run_cuda(sc, norm_dtype=None)
    x = out["x_u8"].cpu().numpy()
    for i in range(12):
        ox, oy, _ = S.run_oracle(sc, i)
        assert np.array_equal(x[i], ox), i
    draws["jit_s"][3] = 2.5
    sc.params = S.ssad_b200.params_from_draws(sc.cfg, draws)
    with pytest.raises(RuntimeError):
        U.run_cuda(sc, norm_dtype=None)
