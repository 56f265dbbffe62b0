"""CPU restatement of the reference's duration alignment (TEST INFRASTRUCTURE ONLY).

Follows Demo/Inference_LJSpeech.ipynb cell 17 and Demo/Inference_LibriTTS.ipynb cell 16 literally: a
one-hot ``pred_aln_trg`` built in a Python loop, ``x @ pred_aln_trg``, and the HiFi-GAN one-frame right
shift.  The reference holds no golden vectors for this step; the restatement is pinned on the notebook's
own statements, cut out of the notebook JSON and executed by oracle/make_golden_alignment.py
(tests/golden/alignment_notebook.npz, tests/test_oracle_golden.py::test_alignment_oracle_vs_notebook_cell)."""
import torch


def one_hot_alignment(dur: torch.Tensor, frames: int) -> torch.Tensor:
    """pred_aln_trg [Ltext, frames] for one item (columns past sum(dur) stay zero)."""
    aln = torch.zeros(dur.shape[0], frames)
    c = 0
    for i in range(dur.shape[0]):
        n = int(dur[i])
        aln[i, c:c + n] = 1
        c += n
    return aln


def expand(x: torch.Tensor, dur: torch.Tensor, frames: int, shift: bool) -> torch.Tensor:
    """x [B, C, Ltext], dur [B, Ltext] -> [B, C, frames] (``x @ pred_aln_trg`` per item, then the shift)."""
    out = torch.stack([x[b] @ one_hot_alignment(dur[b], frames)[:, :frames] for b in range(x.shape[0])])
    if shift:
        new = torch.zeros_like(out)
        new[:, :, 0] = out[:, :, 0]
        new[:, :, 1:] = out[:, :, 0:-1]
        out = new
    return out
