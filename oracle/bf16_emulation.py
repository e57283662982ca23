"""CPU emulation of the BF16-operand / fp32-accumulate arithmetic of the tcgen05 path (oracle; tests only).

Same operators as oracle/melgan_oracle.py (reference: melgan_generator.py:29-78, melgan.py:13-38), but with
every tensor-core operand rounded to bfloat16 exactly where the CUDA path rounds it:
  * weights once, at plan creation;
  * every activation when it is stored to HBM in the canonical bf16 layout (the mel, the projection
    output after LeakyReLU, every stack input x, every block's h after LeakyReLU);
  * lrelu(x) of a residual block is computed from the *stored* bf16 x and rounded again (the in-shared-
    memory prologue of the dilated conv) — a double rounding the emulation reproduces.
Accumulation, bias, tanh and PQMF stay fp32.  The CUDA path must match this to accumulation-order noise;
its distance to the fp32 oracle (the 40 dB bar) is then a property of the arithmetic, not of the kernels.
"""
import torch
import torch.nn.functional as F

from .melgan_oracle import ACT_SLOPE, GeneratorSpec, fold_weight_norm, pqmf_filters, pqmf_synthesis


def bf16_round(x):
    return x.to(torch.bfloat16).to(torch.float32)


def generator_inference_bf16(spec: GeneratorSpec, sd, mel, inference_padding=2, taps=None):
    if any(k.endswith("weight_v") for k in sd):
        sd = fold_weight_norm(sd)
    sd = {k: (bf16_round(v.detach().float()) if k.endswith(".weight") else v.detach().float()) for k, v in sd.items()}
    r, lre = bf16_round, (lambda t: F.leaky_relu(t, ACT_SLOPE))
    with torch.no_grad():
        c = mel.detach().float()
        if inference_padding > 0:
            c = F.pad(c, (inference_padding, inference_padding), mode="replicate")
        pp = (spec.proj_kernel - 1) // 2
        x = F.conv1d(F.pad(r(c), (pp, pp), mode="reflect"), sd["layers.1.weight"], sd["layers.1.bias"])
        if taps is not None:
            taps[1] = x
        half = (spec.res_kernel - 1) // 2
        for s, f in enumerate(spec.upsample_factors):
            iu, op = spec.idx_upsample(s), f % 2
            x = F.conv_transpose1d(r(lre(x)), sd[f"layers.{iu}.weight"], sd[f"layers.{iu}.bias"], stride=f,
                                   padding=f // 2 + op, output_padding=op)
            if taps is not None:
                taps[iu] = x
            pre = f"layers.{spec.idx_stack(s)}"
            for b in range(spec.num_res_blocks):
                d = spec.res_kernel ** b
                xr = r(x)
                a = r(lre(xr))
                h = F.conv1d(F.pad(a, (half * d, half * d), mode="reflect"), sd[f"{pre}.blocks.{b}.2.weight"],
                             sd[f"{pre}.blocks.{b}.2.bias"], dilation=d)
                h = F.conv1d(r(lre(h)), sd[f"{pre}.blocks.{b}.4.weight"], sd[f"{pre}.blocks.{b}.4.bias"])
                x = F.conv1d(xr, sd[f"{pre}.shortcuts.{b}.weight"], sd[f"{pre}.shortcuts.{b}.bias"]) + h
            if taps is not None:
                taps[spec.idx_stack(s)] = x
        ih = spec.idx_head()
        x = torch.tanh(F.conv1d(F.pad(r(lre(x)), (pp, pp), mode="reflect"), sd[f"layers.{ih}.weight"],
                                sd[f"layers.{ih}.bias"]))
        if taps is not None:
            taps[ih + 1] = x
        if spec.pqmf:
            H, G, U = pqmf_filters()
            x = pqmf_synthesis(x, G, U)
    return x
