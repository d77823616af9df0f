"""Golden vectors for the Vocos head + inverse STFT from the reference's OWN witnesses of that arithmetic.

The `vocos` package is absent (third-party, unpinned in pyproject.toml:44), but the reference tree carries a second
implementation of the part after the ConvNeXt backbone, used to export the vocoder for its Triton runtime:
  * `runtime/triton_trtllm/scripts/export_vocoder_to_onnx.py:45-59`  class ISTFTHead: Linear -> chunk -> exp ->
    clip(max=1e2) -> cos / sin -> STFT.inverse(real, imag)
  * `runtime/triton_trtllm/scripts/conv_stft.py:52-234`  class STFT (conv-transpose iSTFT, scipy Hann window,
    window-square envelope), imports torch + scipy only.
`conv_stft.py` is imported by path, unmodified; `ISTFTHead` is lifted out of the export script with `ast` (the script
itself imports huggingface_hub / vocos at module level) and executed unmodified.  Its inverse returns T * hop samples
starting at the same offset (n_fft / 2) as vocos' `torch.istft(center=True)` path, which returns (T - 1) * hop: the
first (T - 1) * hop samples are the same quantity and are what the tests compare.

Run in the build container (needs /root/reference):  python oracle/make_golden_vocos.py
  -> tests/golden/vocos_head_istft.npz   (head weight [1026, 16], bias, inputs h [B, T, 16], audio [B, T * 256])
"""
import ast
import importlib.util
import os
import sys

import numpy as np
import torch
import torch.nn as nn

SCRIPTS = "/root/reference/src/f5_tts/runtime/triton_trtllm/scripts"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_reference_head():
    spec = importlib.util.spec_from_file_location("conv_stft", os.path.join(SCRIPTS, "conv_stft.py"))
    conv_stft = importlib.util.module_from_spec(spec)
    sys.modules["conv_stft"] = conv_stft
    spec.loader.exec_module(conv_stft)
    src = os.path.join(SCRIPTS, "export_vocoder_to_onnx.py")
    tree = ast.parse(open(src, encoding="utf-8").read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "ISTFTHead")
    ns = {"nn": nn, "torch": torch, "STFT": conv_stft.STFT}
    exec(compile(ast.Module(body=[cls], type_ignores=[]), src, "exec"), ns)
    return ns["ISTFTHead"], conv_stft.STFT


def main():
    ISTFTHead, STFT = load_reference_head()
    torch.manual_seed(0)
    in_dim, n_fft, hop = 16, 1024, 256
    head = ISTFTHead(n_fft, hop)
    head.out = nn.Linear(in_dim, n_fft + 2)
    with torch.no_grad():
        # log-magnitudes spread over [-6, 6] so that exp() crosses the clip at 1e2 (ln 100 = 4.6) for some bins
        head.out.weight.copy_(torch.randn(n_fft + 2, in_dim) * 0.6)
        head.out.bias.copy_(torch.randn(n_fft + 2) * 0.5)
    out = {"head.out.weight": head.out.weight.detach().numpy(), "head.out.bias": head.out.bias.detach().numpy()}
    # batch 1 only: the witness's envelope division indexes a [1, 1, L] tensor (`coffidx`, conv_stft.py:227-231), so for
    # batch elements >= 1 it silently skips the normalisation -- measured here, 1.2 max-abs off torch.istft -- i.e. the
    # witness is only a witness for the first element of a batch
    for tag, (B, T) in {"a": (1, 24), "b": (1, 5), "c": (1, 2), "d": (1, 57)}.items():
        h = torch.randn(B, T, in_dim)
        with torch.no_grad():
            audio = head(h)
        assert audio.shape == (B, T * hop), audio.shape
        out[f"h_{tag}"] = h.numpy()
        out[f"audio_{tag}"] = audio.numpy()
    # the bare inverse on seeded (real, imag): pins the iSTFT alone, including non-zero imaginary DC / Nyquist bins
    stft = STFT(fft_len=n_fft, win_hop=hop, win_len=n_fft)
    re, im = torch.randn(1, n_fft // 2 + 1, 12), torch.randn(1, n_fft // 2 + 1, 12)
    with torch.no_grad():
        out["inv_re"], out["inv_im"] = re.numpy(), im.numpy()
        out["inv_audio"] = stft.inverse(input1=re, input2=im, input_type="realimag").numpy()
    path = os.path.join(ROOT, "tests", "golden", "vocos_head_istft.npz")
    np.savez_compressed(path, **{k: np.asarray(v, dtype=np.float32) for k, v in out.items()})
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
