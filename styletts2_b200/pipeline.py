"""Host-to-host decoding loop around ``Decoder.forward`` with the copies off the critical path.

The reference's inference call sites (``Demo/Inference_LJSpeech.ipynb`` cell 17) do
``decoder(asr, F0, N, s).squeeze().cpu().numpy()``: upload, forward, download, one after the other on
one stream.  A serving loop decodes batch after batch, so the upload of batch i+1 and the download of
batch i-1 can run on the copy engines while batch i is on the SMs.  ``HostPipeline`` does exactly that
with two copy streams, ``depth`` device-side input slots and CUDA events; the decoder itself is
unchanged (same module call, same stream semantics).
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Sequence

import torch

_KEYS = ("asr", "F0_curve", "N", "s")


class HostPipeline:
    def __init__(self, decoder, device: torch.device, depth: int = 2, item_offset: int = 0):
        """``item_offset``: global index of row 0 of every submitted batch (a rank decoding rows [lo, hi) of a
        sharded batch passes ``lo`` so that the in-kernel noise equals the unsharded batch's)."""
        if depth < 1:
            raise ValueError("depth must be >= 1")
        self.dec, self.dev, self.depth = decoder, torch.device(device), depth
        self.item_offset = int(item_offset)
        self.h2d = torch.cuda.Stream(self.dev)
        self.d2h = torch.cuda.Stream(self.dev)
        self._slots: List[Dict[str, torch.Tensor]] = [dict() for _ in range(depth)]
        self._in_ready = [torch.cuda.Event() for _ in range(depth)]    # upload of the slot finished
        self._in_free = [torch.cuda.Event() for _ in range(depth)]     # forward has consumed the slot
        self._used = [False] * depth
        self._n = 0

    def _slot_tensors(self, slot: int, host: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        cur = self._slots[slot]
        for k in _KEYS:
            t = cur.get(k)
            if t is None or t.shape != host[k].shape:
                cur[k] = torch.empty(host[k].shape, dtype=torch.float32, device=self.dev)
        return cur

    def submit(self, host: Dict[str, torch.Tensor], out: torch.Tensor, **kw) -> None:
        """enqueue one batch: ``host`` holds pinned fp32 ``asr, F0_curve, N, s``; the waveform lands in
        the pinned tensor ``out`` ([B, 1, 300 T]).  Returns immediately; call :meth:`wait` before
        reading ``out``."""
        slot = self._n % self.depth
        self._n += 1
        compute = torch.cuda.current_stream(self.dev)
        with torch.cuda.stream(self.h2d):
            if self._used[slot]:
                self.h2d.wait_event(self._in_free[slot])
            dev_in = self._slot_tensors(slot, host)
            for k in _KEYS:
                dev_in[k].copy_(host[k], non_blocking=True)
            self._in_ready[slot].record(self.h2d)
        self._used[slot] = True
        compute.wait_event(self._in_ready[slot])
        if self.item_offset:
            kw.setdefault("item_offset", self.item_offset)
        wav = self.dec(dev_in["asr"], dev_in["F0_curve"], dev_in["N"], dev_in["s"], **kw)
        self._in_free[slot].record(compute)
        done = torch.cuda.Event()
        done.record(compute)
        with torch.cuda.stream(self.d2h):
            self.d2h.wait_event(done)
            out.copy_(wav, non_blocking=True)
            wav.record_stream(self.d2h)

    def wait(self) -> None:
        """block until every submitted waveform is in host memory"""
        self.d2h.synchronize()

    def decode(self, batches: Iterable[Dict[str, torch.Tensor]], outs: Sequence[torch.Tensor], **kw) -> None:
        for host, out in zip(batches, outs):
            self.submit(host, out, **kw)
        self.wait()
