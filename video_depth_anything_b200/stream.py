"""The streaming single-image mode: `VideoDepthAnything.infere_single_image` (reference video_depth.py:93-327).

The reference encodes ONE new frame per step, keeps the reassembled features of the last `inference_length +
max(keyframe_list) - 1` frames, runs the temporal part of the head over `inference_length` of them (the first frame,
the keyframes and the most recent ones) and predicts depth for the new frame (plus, with `align_each_new_frame`, for
the keyframes, whose earlier outputs then give a per-frame scale/shift fit).  Everything it does on the host -- cv2
pre-processing, fancy-index gathers of the store, the shifting copy of the store, per-frame device->host copies, the
numpy least squares -- runs here on the device through `engine.StreamPlan`; the host only does index arithmetic.
There is one device->host copy: the final stack.
"""
from __future__ import annotations

import numpy as np
import torch

from . import lib as L
from .driver import WindowedDepthDriver, effective_input_size, network_size
from .engine import StreamPlan


def stream_schedule(inference_length: int, keyframe_list):
    """Context rows and alignment positions of the first max(keyframe_list)+1 predicting steps (video_depth.py:139-185).

    Step s predicts frame f = L-1+s from the rows [f-L+1, f) of the store, with row 0 (the first frame of the video)
    always in position 0 and keyframe j in position j+1 unless its row `L - keyframe_list[j]` is inside the range
    anyway; a keyframe that is not, is fetched from `max(row, f - (keyframe_list[j] + L - len(keyframe_list)))`.  The
    alignment positions are where frame 0 and the keyframes sit inside the context.  The reference's arithmetic is kept
    as is, including rows that have not been written yet (they hold zeros)."""
    L_, nk = int(inference_length), len(keyframe_list)
    lag = [kf + L_ - nk for kf in keyframe_list]
    key_row = [L_ - kf if kf < L_ else j + 1 for j, kf in enumerate(keyframe_list)]
    if len(set(key_row)) != nk:
        raise AssertionError(f"Setup leads to duplicates in the keyframes: {key_row}")
    contexts, align_pos = [], []
    for f in range(L_ - 1, L_ + max(keyframe_list)):
        ctx = [0] + list(range(f - L_ + 2, f))
        pos = [0]
        for j, row in enumerate(key_row):
            if row in ctx:
                pos.append(ctx.index(row))
            else:
                pos.append(j + 1)
                ctx[j + 1] = max(row, f - lag[j])
        contexts.append(ctx)
        align_pos.append(pos)
    return contexts, align_pos


class FeatureStore:
    """Logical rows of the reference's feature tensors -> physical ring slots.  The reference shifts its tensors by
    one row (dropping row 1, video_depth.py:291-294) once they are full; here only this table shifts."""

    def __init__(self, rows: int):
        self.slot = list(range(rows))   # logical row -> physical slot; unwritten rows point at still-zero slots
        self.rows = rows

    def write(self, i: int) -> int:
        """Physical slot that receives frame i's features (video_depth.py:281-296)."""
        if i < self.rows:
            return self.slot[i]
        freed = self.slot.pop(1)
        self.slot.append(freed)
        return freed


class StreamingDepthDriver:
    def __init__(self, model, device="cuda", fp32=False, skip_tmp_block=False, enc_batch=8):
        self.model = model
        self.io = WindowedDepthDriver(model, device=device, fp32=fp32, skip_tmp_block=skip_tmp_block)
        self.device = self.io.device
        self.fp32 = fp32
        self.skip_tmp_block = skip_tmp_block
        self.enc_batch = enc_batch   # frames per encoder launch (1 = strictly frame by frame)
        self.lib = L.load()

    def run(self, frames, input_size=518, inference_length=32, keyframe_list=(0, 12), align_each_new_frame=True,
            return_device=False):
        keyframe_list = list(keyframe_list)
        T, max_kf, nk = int(inference_length), max(keyframe_list), len(keyframe_list)
        n = frames.shape[0]
        fh, fw = frames[0].shape[:2]
        nh, nw = network_size(fh, fw, effective_input_size(fh, fw, input_size))
        contexts, align_pos = stream_schedule(T, keyframe_list)
        rows = T + max_kf - 1
        if n < T:
            raise ValueError("need at least one array to stack")  # np.stack([]) in the reference (video_depth.py:325-327)
        npreds = (T, nk + 2) if align_each_new_frame else (1,)
        if align_each_new_frame:
            # the reference indexes the T-1 gathered context frames with these lists (dpt_temporal.py:188-189)
            for lst in [contexts[0]] + align_pos:
                for v in lst:
                    if v >= T - 1:
                        raise IndexError(f"index {v} is out of bounds for dimension 0 with size {T - 1}")
        with torch.cuda.device(self.device), torch.no_grad():
            plan = self.model.stream_plan(T, nh, nw, self.device, not self.fp32, rows, npreds, self.skip_tmp_block,
                                          enc_batch=self.enc_batch)
            for ring in plan.rings:
                ring.zero_()
            return self._run(plan, frames, n, fh, fw, nh, nw, T, max_kf, contexts, align_pos, rows, align_each_new_frame,
                             return_device)

    def _run(self, plan, frames, n, fh, fw, nh, nw, T, max_kf, contexts, align_pos, rows, align, return_device):
        lib, s = self.lib, L.stream_ptr()
        fe = fh * fw
        R, EB = rows, plan.enc_batch  # incoming slots R .. R + EB - 1: frame i's features land in slot R + i % EB
        store = FeatureStore(rows)
        # ---- host: every step's index block, uploaded once
        blocks, meta = [], []
        for i in range(n):
            inc = R + i % EB
            if i < T - 1:
                blocks.append(plan.index_row(commit=[store.write(i)], src=[inc]))
                meta.append(None)
                continue
            k = min(i - (T - 1), len(contexts) - 1)
            ctx = contexts[k]
            slots = [store.slot[r] for r in ctx]
            if align:
                pred = ctx if i == T - 1 else align_pos[k]
                abs_idx = [ctx[p] for p in align_pos[k]]
                if i >= T + max_kf:  # the store has shifted i - (T + max_kf) + 1 times; row 0 never moves
                    abs_idx = [0 if a == 0 else a + (i - (T + max_kf)) + 1 for a in abs_idx]
            else:
                pred, abs_idx = [], []
            blocks.append(plan.index_row(ctx=slots + [inc], pred=[slots[p] for p in pred] + [inc], p3=list(pred) + [T - 1],
                                         commit=[store.write(i)], src=[inc]))
            meta.append((len(pred) + 1, abs_idx))
        table = torch.stack(blocks).to(self.device, non_blocking=True)

        pre = torch.empty((n, 3, nh, nw), dtype=torch.float32, device=self.device)
        n_out = n if align else n - (T - 1)
        out = torch.empty((n_out, fh, fw), dtype=torch.float32, device=self.device)
        stage = torch.empty((max(plan.npreds), fh, fw), dtype=torch.float32, device=self.device)
        st = torch.zeros(2, dtype=torch.float32, device=self.device)
        scratch = torch.zeros(5, dtype=torch.float64, device=self.device)
        uploaded, count = 0, 0
        for i in range(n):
            if i >= uploaded:
                hi = min(n, uploaded + self.io.upload_chunk)
                self.io._upload_preprocess(frames, uploaded, hi, pre, nh, nw)
                uploaded = hi
            plan.load_indices(table[i])
            if i % EB == 0:  # encoder + reassemble for frames i .. i + EB - 1 (a short last batch re-encodes its tail frame)
                hi = min(n, i + EB)
                if hi > uploaded:
                    self.io._upload_preprocess(frames, uploaded, hi, pre, nh, nw)
                    uploaded = hi
                plan.x_in[:hi - i].copy_(pre[i:hi], non_blocking=True)
                if hi - i < EB:
                    plan.x_in[hi - i:].copy_(pre[hi - 1:hi].expand(EB - (hi - i), -1, -1, -1), non_blocking=True)
                plan.run_stage("encode")
            if meta[i] is not None:
                npred, abs_idx = meta[i]
                plan.run_stage("middle")
                plan.run_stage(f"tail{npred}")
                depth = plan.depths[npred]
                if not align or count == 0:
                    # unaligned outputs: the only frame (no alignment) or the whole first context (video_depth.py:304-306,318)
                    L.check(lib.vda_resize_depth(depth.data_ptr(), L.F32, out[count].data_ptr(), npred, nh, nw, fh, fw, s),
                            "vda_resize_depth")
                    count += npred
                else:
                    # scale/shift of this step's keyframe maps onto their earlier outputs, applied to the new frame,
                    # not clamped (video_depth.py:307-315)
                    L.check(lib.vda_resize_depth(depth.data_ptr(), L.F32, stage.data_ptr(), npred, nh, nw, fh, fw, s),
                            "vda_resize_depth")
                    for j, a in enumerate(abs_idx):
                        phase = (1 if j == 0 else 0) | (2 if j == len(abs_idx) - 1 else 0)
                        L.check(lib.vda_scale_shift(stage[j].data_ptr(), out[a].data_ptr(), fe, scratch.data_ptr(),
                                                    st.data_ptr(), phase, s), "vda_scale_shift")
                    L.check(lib.vda_scale_apply(stage[npred - 1].data_ptr(), out[count].data_ptr(), st.data_ptr(), fe, s),
                            "vda_scale_apply")
                    count += 1
            plan.run_stage("commit")
        assert count == n_out, (count, n_out)
        res = out[1:n] if align else out[:n]  # aligned mode: frame 0 only serves the alignment (video_depth.py:324-327)
        if return_device:
            return res
        return self.io._download(res)
