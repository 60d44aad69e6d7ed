"""Host side of the batched generator: owns the resident device state (cached mask bits, donor
bank, normalisation table) and turns (source images, ParamBatch) into one ``ssad_generate`` call.

PyTorch is used for device memory and streams only; all pixel work happens in libssad_gen.so."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np
import torch

from . import _native as N
from .config import CPP, GeneratorConfig
from .params import IMAGE_PARAMS_DTYPE, ParamBatch

DEBUG_DTYPE = np.dtype([
    ("n_coords", "<i4"), ("sim", "<i4"), ("paste", "<i4", (5, 2)), ("contrast_mean", "<i4"),
    ("label_out", "<i4"), ("avg_rgb", "<i4", (3,)), ("cosine", "<f8"), ("reserved", "<i4", (2,)),
], align=True)


def pack_mask_bits(mask: np.ndarray) -> np.ndarray:
    """bool [..., H, W] -> uint32 [..., H, ceil(W/32)], bit (x & 31) of word (x >> 5)."""
    mask = np.asarray(mask, dtype=bool)
    W = mask.shape[-1]
    pad = (-W) % 32
    if pad:
        mask = np.concatenate([mask, np.zeros(mask.shape[:-1] + (pad,), dtype=bool)], axis=-1)
    b = np.packbits(mask, axis=-1, bitorder="little")
    return np.ascontiguousarray(b).view("<u4")


def normalize_lut(dtype: torch.dtype) -> torch.Tensor:
    """[3, 256] table of Normalize(ToTensor(v)) computed with the very torch ops the reference's
    transform runs (datasets.py:430-433): v.float().div(255).sub(mean).div(std)."""
    v = torch.arange(256, dtype=torch.uint8).to(torch.float32).div(255)
    mean = torch.as_tensor(CPP.normalize_mean, dtype=torch.float32)[:, None]
    std = torch.as_tensor(CPP.normalize_std, dtype=torch.float32)[:, None]
    lut = (v[None, :] - mean) / std
    return lut.to(dtype).contiguous()


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


class Plan:
    """Frozen arguments of one ssad_generate call (keeps the tensors alive)."""


class BatchGenerator:
    """Batched replacement of ``PretextTaskDataset.__getitem__`` (reference datasets.py:209-394).

    mask:   None for textures (all-white, datasets.py:197-198); bool [H, W] for fixed-pose classes (the
            cached ``obj_mask`` of image 000, datasets.py:199-206); per-image masks are passed to
            ``generate`` (non-fixed classes, datasets.py:232-233).
    donors: uint8 [D, H, W, 3] bank of ``images_for_cut`` (datasets.py:189-193), textures only.
    """

    def __init__(self, cfg: GeneratorConfig, device="cuda", mask: Optional[np.ndarray] = None,
                 donors: Optional[np.ndarray] = None):
        self.cfg = cfg
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("BatchGenerator needs a CUDA device: there is no CPU fallback")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.lib = N.load()
        W, H = cfg.imsize
        self.mask_bits = None
        self.mask_host = None
        if mask is not None:
            mask = np.asarray(mask, dtype=bool)
            self.mask_host = mask
            if mask.shape != (H, W):
                raise ValueError(f"mask must be [H={H}, W={W}]")
            self.mask_bits = torch.from_numpy(pack_mask_bits(mask).view(np.int32)).to(self.device)
        self.donors = None
        if donors is not None:
            donors = np.ascontiguousarray(donors, dtype=np.uint8)
            if donors.ndim != 4 or donors.shape[1:] != (H, W, 3):
                raise ValueError("donors must be uint8 [D, H, W, 3] at imsize")
            self.donors = torch.from_numpy(donors).to(self.device)
        if cfg.is_texture and self.donors is None:
            raise ValueError("texture classes need the donor bank (datasets.py:189-193,226)")
        self._luts = {}

    def lut(self, dtype: torch.dtype) -> torch.Tensor:
        if dtype not in self._luts:
            self._luts[dtype] = normalize_lut(dtype).to(self.device)
        return self._luts[dtype]

    # ------------------------------------------------------------------
    def validate_records(self, rec: np.ndarray, src_shape, n_src: Optional[int] = None) -> None:
        """Host-side check of every record field the kernels index memory with (the library itself trusts the records):
        raises ValueError instead of letting a bad record read out of bounds on the device."""
        cfg = self.cfg
        if rec.dtype != IMAGE_PARAMS_DTYPE:
            raise ValueError("records have the wrong dtype")
        if len(rec) == 0:
            return
        W, H = cfg.work_size
        sh, sw = int(src_shape[0]), int(src_shape[1])
        if np.any(rec["jit_f"] < 0) or np.any(rec["bright_f"] < 0):
            raise ValueError("enhancement factors must be non-negative")
        if int(rec["label"].max()) > 3:
            raise ValueError("labels must be in {0, 1, 2, 3}")
        if int(rec["poly_n"].max()) > 8:
            raise ValueError("a polygon has at most 8 vertices")
        nd = 0 if self.donors is None else int(self.donors.shape[0])
        dmin, dmax = int(rec["donor_idx"].min()), int(rec["donor_idx"].max())
        if dmin < -1 or (dmax >= 0 and dmax >= nd):
            raise ValueError("donor_idx outside the donor bank")
        co = rec["crop_off"].astype(np.int64)
        if co.min() < 0 or (co[:, 0] + W).max() > sw or (co[:, 1] + H).max() > sh:
            raise ValueError("crop window outside the source image")
        do = rec["donor_off"].astype(np.int64)
        dW, dH = cfg.donor_window
        if do.min() < 0 or (do[:, 0] + dW).max() > sw or (do[:, 1] + dH).max() > sh:
            raise ValueError("donor window outside the donor image")
        r = rec["rect"].astype(np.int64)
        if r[:, :2].min() < 0 or r[:, 2:].min() < 0:
            raise ValueError("patch rectangle must be non-negative")
        scar = rec["label"] == 2
        if np.any(scar) and (int(rec["scar_k"][scar].min()) < 1 or int(rec["scar_k"][scar].max()) > 5):
            raise ValueError("scar_k must be in 1..5")
        if n_src is not None and int(rec["src_index"].max()) >= n_src:
            raise ValueError("src_index outside the resident set")

    def upload_params(self, params: ParamBatch, src_index=None) -> torch.Tensor:
        rec = np.ascontiguousarray(params.records)
        if rec.dtype != IMAGE_PARAMS_DTYPE:
            raise ValueError("records have the wrong dtype")
        if src_index is not None:
            rec = rec.copy()
            rec["src_index"] = np.asarray(src_index, dtype=np.uint32)
        t = torch.from_numpy(rec.view(np.uint8).reshape(len(rec), IMAGE_PARAMS_DTYPE.itemsize))
        return t.to(self.device, non_blocking=True)

    def describe(self, batch: int, src: torch.Tensor, max_pw: int, max_ph: int, masks, norm_dtype, crop_aligned=False,
                 indexed: bool = False) -> N.SsadBatchDesc:
        cfg = self.cfg
        W, H = cfg.work_size
        d = N.SsadBatchDesc()
        d.batch = batch
        d.height, d.width = H, W
        d.src_height, d.src_width = int(src.shape[1]), int(src.shape[2])
        d.src_pitch, d.src_stride = int(src.stride(1)), int(src.stride(0))
        if self.donors is not None:
            d.donor_count = int(self.donors.shape[0])
            d.donor_height, d.donor_width = int(self.donors.shape[1]), int(self.donors.shape[2])
            d.donor_pitch, d.donor_stride = int(self.donors.stride(1)), int(self.donors.stride(0))
        d.donor_win_w, d.donor_win_h = cfg.donor_window
        if masks is not None:
            d.mask_mode = N.SSAD_MASK_PER_IMAGE
            d.mask_pitch_words, d.mask_stride_words = int(masks.stride(1)), int(masks.stride(0))
            d.mask_height, d.mask_width = int(masks.shape[1]), cfg.imsize[0]
        elif self.mask_bits is not None:
            d.mask_mode = N.SSAD_MASK_SHARED
            d.mask_pitch_words = int(self.mask_bits.stride(0))
            d.mask_height, d.mask_width = int(self.mask_bits.shape[0]), cfg.imsize[0]
        else:
            d.mask_mode = N.SSAD_MASK_ALL
        d.box_poly = (C.c_int32 * 4)(*cfg.box_poly)
        d.box_scar = (C.c_int32 * 4)(*cfg.box_scar)
        d.coverage_min3 = cfg.coverage_min3
        d.max_patch_w, d.max_patch_h = max(int(max_pw), 1), max(int(max_ph), 1)
        d.out_norm_dtype = {None: N.SSAD_NORM_NONE, torch.float32: N.SSAD_NORM_F32, torch.bfloat16: N.SSAD_NORM_BF16}[norm_dtype]
        d.out_u8_pitch, d.out_u8_stride = 3 * W, 3 * W * H
        d.crop_align16 = int(crop_aligned)
        d.any_warp = int(cfg.warp)
        d.src_indexed = int(indexed)
        d.src_count = int(src.shape[0]) if indexed else 0
        d.line_width = 1 if cfg.line_patch_mode else 3             # datasets.py:383-388
        return d

    def _window_mask_fn(self, rec: np.ndarray, masks, src_index):
        """i -> bool [h, w] mask window of item i as ``__getitem__`` sees it (datasets.py:242-264), None for textures."""
        W, H = self.cfg.work_size
        host_masks = None
        if masks is not None:                        # per-image masks live on the device: label-3 items need theirs here
            host_masks = unpack_mask_bits(masks.cpu().numpy(), self.cfg.imsize[0])

        def fn(i: int):
            if host_masks is not None:
                m = host_masks[int(src_index[i]) if src_index is not None else i]
            elif self.mask_host is not None:
                m = self.mask_host
            else:
                return None
            x0, y0 = int(rec["crop_off"][i][0]), int(rec["crop_off"][i][1])
            return m[y0:y0 + H, x0:x0 + W]
        return fn

    @staticmethod
    def batch_hints(rec: np.ndarray):
        """Batch-wide bounds the descriptor carries: (crop offsets 16-byte aligned, canvas width, canvas height, largest
        enabled Saturation factor).  The polygon lives on pw x ph, the rotated scar on scar_size (datasets.py:344)."""
        crop_aligned = bool(np.all((rec["crop_off"][:, 0].astype(np.int64) * 3) % 16 == 0))
        max_w = max(int(rec["rect"][:, 2].max(initial=1)), int(rec["scar_size"][:, 0].max(initial=1)))
        max_h = max(int(rec["rect"][:, 3].max(initial=1)), int(rec["scar_size"][:, 1].max(initial=1)))
        sat_on = (rec["jit_en"] & 4) != 0
        scar = (max(int(rec["scar_size"][:, 0].max(initial=1)), 1), max(int(rec["scar_size"][:, 1].max(initial=1)), 1))
        return (crop_aligned, max_w, max_h, float(rec["jit_f"][:, 2][sat_on].max(initial=0.0)), scar,
                bool(np.any(rec["label"] == 3)))

    def plan(self, src: torch.Tensor, params: ParamBatch, *, masks=None, want_u8=True, norm_dtype=None, debug=False,
             params_dev=None, out_u8=None, out_norm=None, y_out=None, workspace=None, src_index=None,
             context=None, validated: bool = False, hints=None) -> "Plan":
        """Validate once and freeze descriptor + device pointers of one ssad_generate call; ``run(plan)``
        then costs one FFI call (steady-state loops, CUDA-graph capture).  ``src_index``: item i works on
        ``src[src_index[i]]`` (``src`` = the resident set, never gathered); per-image ``masks`` are then per source image.
        ``context``: a handle from ``new_context()`` for a host thread that shares the device with other threads."""
        cfg = self.cfg
        B = len(params)
        W, H = cfg.work_size
        rec = params.records
        indexed = src_index is not None
        if not indexed and int(src.shape[0]) != B:
            raise ValueError("params and src disagree on the batch size")
        if indexed and len(src_index) != B:
            raise ValueError("params and src_index disagree on the batch size")
        if not validated:
            self.validate_records(rec, src.shape[1:3])
        if indexed and (int(np.min(src_index)) < 0 or int(np.max(src_index)) >= int(src.shape[0])):
            raise ValueError("src_index outside the resident set")
        params_dev_given = params_dev is not None
        if params_dev is None:
            params_dev = self.upload_params(params, src_index)
        if hints is None:
            hints = self.batch_hints(rec)
        crop_aligned, max_w, max_h, sat_max, scar_dims, any_line = hints
        line_dev = None
        if any_line:
            # label 3 "long line": the vertices depend on the mask (datasets.py:357-377) -- resolved on the host once
            if params.line_xy is None:
                if params_dev_given:
                    raise ValueError("label-3 records: call params.resolve_lines(...) before uploading the records")
                params.resolve_lines(self._window_mask_fn(rec, masks, src_index))
            max_w, max_h = max(max_w, W), max(max_h, H)      # the coverage rows span the poly-line's bounding box
        d = self.describe(B, src, max_w, max_h, masks, norm_dtype, crop_aligned, indexed)
        d.max_scar_w, d.max_scar_h = scar_dims
        if any_line:
            line_dev = torch.from_numpy(np.ascontiguousarray(params.line_xy, dtype=np.int16)).to(self.device)
            d.line_max = int(params.line_xy.shape[1])
            d.line_xy = line_dev.data_ptr()
            if params_dev_given is False:
                params_dev = self.upload_params(params, src_index)   # (line_n was written by resolve_lines)
        d.sat_factor_max = sat_max
        if want_u8 and out_u8 is None:
            out_u8 = torch.empty((B, H, W, 3), dtype=torch.uint8, device=self.device)
        if norm_dtype is not None and out_norm is None:
            out_norm = torch.empty((B, 3, H, W), dtype=norm_dtype, device=self.device)
        if y_out is None:
            y_out = torch.empty((B,), dtype=torch.int64, device=self.device)
        dbg = torch.zeros((B, DEBUG_DTYPE.itemsize), dtype=torch.uint8, device=self.device) if debug else None
        mask_t = masks if masks is not None else self.mask_bits
        lut = self.lut(norm_dtype) if norm_dtype is not None else None
        need = int(self.lib.ssad_generate_workspace(C.byref(d)))
        if workspace is None or workspace.numel() * workspace.element_size() < need:
            workspace = torch.empty((max(need, 256),), dtype=torch.uint8, device=self.device)
        pl = Plan()
        pl.line_dev, pl.context = line_dev, context
        pl.workspace = workspace
        pl.desc, pl.params_dev, pl.src, pl.mask_t, pl.lut = d, params_dev, src, mask_t, lut
        pl.out_u8, pl.out_norm, pl.y_out, pl.dbg = (out_u8 if want_u8 else None), out_norm, y_out, dbg
        pl.args = (C.byref(d), _ptr(params_dev), _ptr(src), _ptr(self.donors), _ptr(mask_t), _ptr(lut),
                   _ptr(pl.out_u8), _ptr(out_norm), _ptr(y_out), _ptr(dbg), _ptr(workspace),
                   workspace.numel() * workspace.element_size())
        return pl

    def new_context(self):
        """A library context (its own helper streams) for a host thread that drives this device next to other threads."""
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            N.check(self.lib.ssad_context_create(C.byref(h)), "ssad_context_create")
        return h

    def free_context(self, ctx) -> None:
        self.lib.ssad_context_destroy(ctx)

    def _call(self, plan: "Plan") -> None:
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream(self.device).cuda_stream
            N.check(self.lib.ssad_generate_ctx(plan.context, *plan.args, C.c_void_p(stream)), "ssad_generate")

    def run(self, plan: "Plan") -> None:
        graph = getattr(plan, "graph", None)
        if graph is not None:
            graph.replay()
            return
        self._call(plan)

    def capture(self, plan: "Plan") -> "Plan":
        """Record the plan's ssad_generate call (a handful of kernels, forked over helper streams) into a CUDA graph:
        ``run(plan)`` then replays it with one launch.  The library never allocates or synchronises, so the call is
        capturable as it is; inputs are read from, and outputs written to, the plan's tensors on every replay."""
        with torch.cuda.device(self.device):
            self.run(plan)                       # warm-up outside the capture (lazy module load, helper streams)
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._call(plan)
            plan.graph = g
        return plan

    def generate(self, src: torch.Tensor, params: ParamBatch, *, masks: Optional[torch.Tensor] = None,
                 want_u8: bool = True, norm_dtype: Optional[torch.dtype] = None, debug: bool = False,
                 params_dev: Optional[torch.Tensor] = None, out_u8: Optional[torch.Tensor] = None,
                 out_norm: Optional[torch.Tensor] = None, y_out: Optional[torch.Tensor] = None,
                 src_index=None, context=None):
        """src: uint8 [B, Hs, Ws, 3] on the device (the decoded, resized originals); with ``src_index`` the resident set
        [N, Hs, Ws, 3] that item i indexes as ``src[src_index[i]]``.
        masks: int32/uint32-packed [B, Hs, ceil(Ws/32)] per-image mask bits (non-fixed classes).
        Returns dict(x_u8, x, y, debug) -- entries present per the want_* flags."""
        cfg = self.cfg
        if src.device != self.device or src.dtype != torch.uint8 or src.dim() != 4 or src.shape[-1] != 3:
            raise ValueError("src must be a uint8 [B, H, W, 3] tensor on the generator's device")
        if src.stride(3) != 1 or src.stride(2) != 3:
            raise ValueError("src pixels must be interleaved RGB with unit channel stride")
        B = len(params)
        if src_index is None and int(src.shape[0]) != B:
            raise ValueError("params and src disagree on the batch size")
        iW, iH = cfg.imsize
        if (int(src.shape[1]), int(src.shape[2])) != (iH, iW):
            raise ValueError(f"src images must be at imsize {cfg.imsize}")
        if cfg.is_non_fixed and masks is None and self.mask_bits is None:
            raise ValueError("non-fixed classes need per-image masks (obj_mask, datasets.py:232-233)")
        plan = self.plan(src, params, masks=masks, want_u8=want_u8, norm_dtype=norm_dtype, debug=debug,
                         params_dev=params_dev, out_u8=out_u8, out_norm=out_norm, y_out=y_out, src_index=src_index,
                         context=context)
        self.run(plan)
        out_u8, out_norm, y_out, dbg = plan.out_u8, plan.out_norm, plan.y_out, plan.dbg
        out = {"y": y_out}
        if want_u8:
            out["x_u8"] = out_u8
        if norm_dtype is not None:
            out["x"] = out_norm
        if debug:
            out["debug"] = dbg.cpu().numpy().view(DEBUG_DTYPE).reshape(B)
        return out


    # ------------------------------------------------------------------ host buffers in, host buffers out
    def generate_from_host(self, h_src: torch.Tensor, params: ParamBatch, h_out: torch.Tensor, h_y: torch.Tensor, *,
                           chunks: Optional[int] = None, masks_fn=None, wait: bool = True, upload: str = "auto"):
        """The end-to-end call for data that lives in (pinned) host memory: uint8 images ``h_src`` [B, Hs, Ws, 3] ->
        ``h_out`` uint8 [B, h, w, 3] and labels ``h_y`` int64 [B], all host tensors.  The batch is cut into ``chunks``
        pieces that flow through three streams -- host-to-device copy, generation (``ssad_generate``), device-to-host
        copy -- so the two PCIe directions and the kernels overlap.  Returns when the results are in host memory;
        with ``wait=False`` it returns at once with a ``torch.cuda.Event`` that completes when they are (a loader that
        keeps the next batch in flight: call again with other host buffers, then ``event.synchronize()`` this one).
        ``masks_fn(d_src_chunk) -> packed masks`` supplies per-image masks (non-fixed classes: ``obj_mask_batch``).
        ``upload``: "full" (= "auto") copies every image whole with the copy engine; "windows" fetches only what a
        patch-mode call reads (the working window and the cut rectangle, ``ssad_upload_windows``: a kernel reads the pinned
        host memory directly) -- 36 % fewer bytes at BASELINE config 2 and 2.30 ms instead of 3.15 ms for the upload on
        its own, but with the results travelling the other way at the same time the kernel's reads drop to 30 GB/s and the
        pair takes as long as the whole-image copy (3.29 vs 3.25 ms, tools/time_upload.py), so it is opt-in: for callers
        whose results stay on the device.  Needs patch mode, no pre-warp, no per-image ``obj_mask``."""
        B = int(h_src.shape[0])
        W, H = self.cfg.work_size
        if chunks is None:
            chunks = min(8, B // 64)             # pieces of at least 64 images, eight at most
        chunks = max(1, min(int(chunks), B))
        st = self._host_pipe(B, tuple(h_src.shape[1:]), chunks)
        cur = torch.cuda.current_stream(self.device)
        for s_ in (st["up"], st["run"], st["down"]):
            s_.wait_stream(cur)
        # the device staging buffers are shared by consecutive calls: this call's uploads follow the previous call's
        # kernels, its kernels the previous call's downloads (both long finished unless the call was made with wait=False)
        st["up"].wait_stream(st["run"])
        st["run"].wait_stream(st["down"])
        bounds = [(B * c // chunks, B * (c + 1) // chunks) for c in range(chunks)]
        # the records of the whole batch go up once, through a pinned staging buffer, ahead of the first pixels
        rec = np.ascontiguousarray(params.records)
        if len(rec) != B or rec.dtype != IMAGE_PARAMS_DTYPE:
            raise ValueError("params and h_src disagree on the batch size (or the records have the wrong dtype)")
        self.validate_records(rec, h_src.shape[1:3])
        hints = self.batch_hints(rec)            # one set of bounds for every piece of the batch
        if hints[5] and params.line_xy is None:  # label 3: the poly-lines are resolved against the mask before the upload
            if masks_fn is not None:
                raise ValueError("label-3 records with per-image device masks: call params.resolve_lines(...) first")
            params.resolve_lines(self._window_mask_fn(rec, None, None))
            rec = np.ascontiguousarray(params.records)
        Hs, Ws = int(h_src.shape[1]), int(h_src.shape[2])
        win_ok = (masks_fn is None and self.cfg.patch_localization and not rec["warp_en"].any() and h_src.is_pinned()
                  and h_src.is_contiguous() and h_src.data_ptr() % 16 == 0 and (Hs * Ws * 3) % 16 == 0)
        if upload == "windows" and not win_ok:
            raise ValueError("upload='windows' needs patch mode, pinned contiguous 16-byte aligned host images, no pre-warp and no masks_fn")
        if upload not in ("auto", "full", "windows"):
            raise ValueError("upload must be 'auto', 'full' or 'windows'")
        use_win = upload == "windows"
        st["ev_rec"].synchronize()               # the previous call's upload has left the staging buffer
        st["h_rec"].numpy()[...] = rec.view(np.uint8).reshape(B, IMAGE_PARAMS_DTYPE.itemsize)
        with torch.cuda.stream(st["up"]):
            st["d_rec"].copy_(st["h_rec"], non_blocking=True)
            st["ev_rec"].record(st["up"])
        for c, (lo, hi) in enumerate(bounds):
            d_src, d_out, d_y = st["src"][lo:hi], st["out"][lo:hi], st["y"][lo:hi]
            with torch.cuda.stream(st["up"]):
                if use_win:
                    N.check(self.lib.ssad_upload_windows(hi - lo, C.c_void_p(st["d_rec"][lo:hi].data_ptr()),
                                                         C.c_void_p(h_src[lo:hi].data_ptr()), C.c_void_p(d_src.data_ptr()),
                                                         Ws * 3, Hs * Ws * 3, Hs, Ws, H, W, int(hints[2]),
                                                         C.c_void_p(st["up"].cuda_stream)), "ssad_upload_windows")
                else:
                    d_src.copy_(h_src[lo:hi], non_blocking=True)
                st["ev_up"][c].record(st["up"])
            with torch.cuda.stream(st["run"]):
                st["run"].wait_event(st["ev_up"][c])
                m = masks_fn(d_src) if masks_fn is not None else None
                if self.cfg.is_non_fixed and m is None and self.mask_bits is None:
                    raise ValueError("non-fixed classes need per-image masks (obj_mask, datasets.py:232-233)")
                pl = self.plan(d_src, params.slice(lo, hi), masks=m, want_u8=True, norm_dtype=None,
                               params_dev=st["d_rec"][lo:hi], out_u8=d_out, y_out=d_y, workspace=st["ws"][c], validated=True,
                               hints=hints)
                st["ws"][c] = pl.workspace       # kept: the next call's chunk reuses it
                self.run(pl)
                st["ev_run"][c].record(st["run"])
            with torch.cuda.stream(st["down"]):
                st["down"].wait_event(st["ev_run"][c])
                h_out[lo:hi].copy_(d_out, non_blocking=True)
                h_y[lo:hi].copy_(d_y, non_blocking=True)
        done = torch.cuda.Event()
        done.record(st["down"])
        cur.wait_stream(st["up"]); cur.wait_stream(st["run"])
        if wait:
            done.synchronize()                   # the caller owns the results when the call returns
        return done

    def generate_resident_to_host(self, src_set: torch.Tensor, src_index, params: ParamBatch, h_out: torch.Tensor,
                                  h_y: torch.Tensor, *, masks: Optional[torch.Tensor] = None, chunks: Optional[int] = None,
                                  wait: bool = True):
        """The deployment the reference's DataModule has (datasets.py:476-508: the decoded set is loaded once): the
        distinct source images ``src_set`` [N, Hs, Ws, 3] stay on the device, item i works on ``src_set[src_index[i]]``;
        per step only the 192-byte records go up and the results ``h_out`` uint8 [B, h, w, 3] / ``h_y`` int64 [B] (pinned
        host tensors) come down, pieces of the batch pipelined through a generation and a download stream.  ``masks``: per
        SOURCE image (non-fixed classes: ``obj_mask_batch(src_set)``, computed once -- obj_mask depends on the original
        only).  Returns a ``torch.cuda.Event`` (complete when the results are in host memory; ``wait=True`` waits)."""
        B = len(params)
        W, H = self.cfg.work_size
        if chunks is None:
            chunks = min(8, B // 64)
        chunks = max(1, min(int(chunks), B))
        st = self._host_pipe(B, (1, 1, 3), chunks)           # (no device staging of sources: a 1x1 placeholder)
        src_index = np.asarray(src_index, dtype=np.int64)
        if len(src_index) != B:
            raise ValueError("params and src_index disagree on the batch size")
        if self.cfg.is_non_fixed and masks is None and self.mask_bits is None:
            raise ValueError("non-fixed classes need per-image masks (obj_mask, datasets.py:232-233)")
        cur = torch.cuda.current_stream(self.device)
        for s_ in (st["up"], st["run"], st["down"]):
            s_.wait_stream(cur)
        st["run"].wait_stream(st["down"])            # the staging outputs are shared with the previous call's downloads
        rec = np.ascontiguousarray(params.records)
        if rec.dtype != IMAGE_PARAMS_DTYPE:
            raise ValueError("records have the wrong dtype")
        rec = rec.copy()
        rec["src_index"] = src_index.astype(np.uint32)
        self.validate_records(rec, src_set.shape[1:3], n_src=int(src_set.shape[0]))
        hints = self.batch_hints(rec)
        if hints[5] and params.line_xy is None:
            params.resolve_lines(self._window_mask_fn(rec, masks, src_index))
            rec["line_n"] = params.records["line_n"]
        st["ev_rec"].synchronize()
        st["h_rec"].numpy()[...] = rec.view(np.uint8).reshape(B, IMAGE_PARAMS_DTYPE.itemsize)
        with torch.cuda.stream(st["up"]):
            st["d_rec"].copy_(st["h_rec"], non_blocking=True)
            st["ev_rec"].record(st["up"])
        st["run"].wait_event(st["ev_rec"])
        bounds = [(B * c // chunks, B * (c + 1) // chunks) for c in range(chunks)]
        for c, (lo, hi) in enumerate(bounds):
            d_out, d_y = st["out"][lo:hi], st["y"][lo:hi]
            with torch.cuda.stream(st["run"]):
                pl = self.plan(src_set, params.slice(lo, hi), masks=masks, want_u8=True, norm_dtype=None,
                               params_dev=st["d_rec"][lo:hi], out_u8=d_out, y_out=d_y, workspace=st["ws"][c],
                               src_index=src_index[lo:hi], validated=True, hints=hints)
                st["ws"][c] = pl.workspace
                self.run(pl)
                st["ev_run"][c].record(st["run"])
            with torch.cuda.stream(st["down"]):
                st["down"].wait_event(st["ev_run"][c])
                h_out[lo:hi].copy_(d_out, non_blocking=True)
                h_y[lo:hi].copy_(d_y, non_blocking=True)
        done = torch.cuda.Event()
        done.record(st["down"])
        cur.wait_stream(st["up"]); cur.wait_stream(st["run"])
        if wait:
            done.synchronize()
        return done

    def _host_pipe(self, B: int, src_shape: tuple, chunks: int) -> dict:
        key = (B, src_shape, chunks)
        st = getattr(self, "_pipe", None)
        if st is None or st["key"] != key:
            W, H = self.cfg.work_size
            dev = self.device
            st = {"key": key, "up": torch.cuda.Stream(dev), "run": torch.cuda.Stream(dev), "down": torch.cuda.Stream(dev),
                  "src": torch.empty((B,) + src_shape, dtype=torch.uint8, device=dev),
                  "out": torch.empty((B, H, W, 3), dtype=torch.uint8, device=dev),
                  "y": torch.empty((B,), dtype=torch.int64, device=dev),
                  "h_rec": torch.empty((B, IMAGE_PARAMS_DTYPE.itemsize), dtype=torch.uint8).pin_memory(),
                  "d_rec": torch.empty((B, IMAGE_PARAMS_DTYPE.itemsize), dtype=torch.uint8, device=dev),
                  "ev_rec": torch.cuda.Event(), "ws": [None] * chunks,
                  "ev_up": [torch.cuda.Event() for _ in range(chunks)], "ev_run": [torch.cuda.Event() for _ in range(chunks)]}
            self._pipe = st
        return st


# ---------------------------------------------------------------------- obj_mask on device
def canny_consts(sigma: float = 1.5, low_threshold: float = 5, high_threshold: float = 15) -> "N.SsadCannyConsts":
    """Constants of ``canny(gray, sigma=1.5, low_threshold=5, high_threshold=15)`` (dataset_generator.py:30) for a
    uint8 image, computed as the CPU libraries compute them: taps as ``scipy.ndimage._filters._gaussian_kernel1d``
    (truncate 4.0 -> radius 6), thresholds divided by the dtype max."""
    radius = int(4.0 * float(sigma) + 0.5)
    if radius != 6:
        raise ValueError("the device kernel is built for radius 6 (sigma = 1.5)")
    sigma2 = sigma * sigma
    x = np.arange(-radius, radius + 1)
    phi = np.exp(-0.5 / sigma2 * x ** 2)
    phi = (phi / phi.sum())[::-1]
    k = N.SsadCannyConsts()
    k.gauss_w = (C.c_double * 13)(*[float(v) for v in phi])
    k.low, k.high = low_threshold / 255.0, high_threshold / 255.0
    k.eps = float(np.finfo(np.float64).eps)
    return k


def unpack_mask_bits(bits: np.ndarray, width: int) -> np.ndarray:
    """uint32/int32 [..., H, nw] -> bool [..., H, width] (inverse of ``pack_mask_bits``)."""
    b = np.ascontiguousarray(bits).view(np.uint8)
    return np.unpackbits(b, axis=-1, bitorder="little")[..., :width].astype(bool)


def obj_mask_batch(src: torch.Tensor, out: Optional[torch.Tensor] = None, workspace: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``obj_mask`` (dataset_generator.py:27-39) for every image of ``src`` (uint8 [B, H, W, 3] on a CUDA device)
    -> bit-packed masks int32 [B, H, ceil(W/32)] on the same device (the layout ``BatchGenerator.generate(masks=...)``
    takes).  Runs on the current stream; no host synchronisation."""
    if src.device.type != "cuda" or src.dtype != torch.uint8 or src.dim() != 4 or src.shape[-1] != 3:
        raise ValueError("src must be a uint8 [B, H, W, 3] CUDA tensor")
    if src.stride(3) != 1 or src.stride(2) != 3:
        raise ValueError("src pixels must be interleaved RGB with unit channel stride")
    lib = N.load()
    B, H, W = int(src.shape[0]), int(src.shape[1]), int(src.shape[2])
    nw = (W + 31) // 32
    if out is None:
        out = torch.empty((B, H, nw), dtype=torch.int32, device=src.device)
    need = int(lib.ssad_obj_mask_workspace(B, H, W))
    if workspace is None or workspace.numel() * workspace.element_size() < need:
        workspace = torch.empty((max(need, 16),), dtype=torch.uint8, device=src.device)
    k = canny_consts()
    with torch.cuda.device(src.device):
        stream = torch.cuda.current_stream(src.device).cuda_stream
        N.check(lib.ssad_obj_mask(B, H, W, C.c_void_p(src.data_ptr()), int(src.stride(1)), int(src.stride(0)), C.byref(k),
                                  C.c_void_p(out.data_ptr()), int(out.stride(1)), int(out.stride(0)),
                                  C.c_void_p(workspace.data_ptr()), workspace.numel() * workspace.element_size(),
                                  C.c_void_p(stream)), "ssad_obj_mask")
    return out
