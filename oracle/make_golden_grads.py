"""Golden GRADIENTS of the flow-matching loss from the UNMODIFIED reference: `CFM.forward` (cfm.py:231-302) under the
seeded draws of oracle/make_golden.py's loss cases, then `loss.backward()` (what trainer.py:384 does), for the
parameters the B200 backward pass covers: the last DiTBlock, norm_out, proj_out and the residual stream entering that
block in full; every other parameter (earlier blocks, timestep MLP, input_embed, text_embed) and the residual stream entering block 0
as every 5th element of the flattened gradient (captured with forward pre-hooks).

Run in the build container (needs /root/reference):  python oracle/make_golden_grads.py
  -> tests/golden/grads_<case>.npz
"""
import os
import random
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import f5_oracle as O, ref_loader  # noqa: E402
from oracle.make_golden import build_reference, case_inputs  # noqa: E402

GRAD_CASES = (("dit_v1style_b3_mask", 21), ("dit_smallstyle_b2_nomask", 23))
SUB = 5


def tail_names(depth):
    blk = f"transformer.transformer_blocks.{depth - 1}."
    return lambda n: n.startswith(blk) or n.startswith("transformer.norm_out.") or n.startswith("transformer.proj_out.")


def main():
    R = ref_loader.load()
    out_dir = os.path.join(ROOT, "tests", "golden")
    for name, seed in GRAD_CASES:
        cfg, cond, text, lens, *_ = case_inputs(name)
        sd = O.synth_backbone_weights(cfg, seed=11)
        cfm = build_reference(R, cfg, sd).train(False)
        x1 = cond[:, : int(lens.max())]
        captured = {}

        def hook(mod, args, kwargs):
            x = args[0] if args else kwargs["x"]
            x.retain_grad()
            captured["x"] = x

        def hook0(mod, args, kwargs):
            x = args[0] if args else kwargs["x"]
            x.retain_grad()
            captured["x0"] = x

        h = cfm.transformer.transformer_blocks[-1].register_forward_pre_hook(hook, with_kwargs=True)
        h0 = cfm.transformer.transformer_blocks[0].register_forward_pre_hook(hook0, with_kwargs=True)
        torch.manual_seed(seed)
        random.seed(seed)
        loss, _, pred = cfm(x1, text, lens=lens)
        loss.backward()
        h.remove()
        h0.remove()
        keep = tail_names(cfg.depth)
        out = {n: p.grad.numpy() for n, p in cfm.named_parameters() if keep(n) and p.grad is not None}
        # the earlier blocks and the timestep MLP: every SUB-th element of the flattened gradient (keeps the fixture small)
        for n, p in cfm.named_parameters():
            if not keep(n) and p.grad is not None and n.startswith("transformer."):
                out[f"sub{SUB}:" + n] = p.grad.flatten()[::SUB].numpy()
        out["d_block_input"] = captured["x"].grad.numpy()
        out[f"sub{SUB}:d_input_embed"] = captured["x0"].grad.flatten()[::SUB].numpy()
        out["loss"] = loss.detach().numpy()
        out["seed"] = np.int64(seed)
        path = os.path.join(out_dir, f"grads_{name}.npz")
        np.savez_compressed(path, **out)
        print(name, "loss", float(loss.detach()), len(out) - 2, "gradients,", os.path.getsize(path), "bytes",
              "max |grad|", max(float(np.abs(v).max()) for k, v in out.items() if k not in ("loss", "seed")))


if __name__ == "__main__":
    main()
