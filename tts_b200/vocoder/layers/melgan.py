"""ResidualStack with the reference's parameter layout (TTS/vocoder/layers/melgan.py:5-45).

This module is the *parameter container*: it keeps the reference's sub-module names so that
checkpoints load with strict keys (``blocks.B.2.weight_g`` ..., ``shortcuts.B.*``).  Inference
of a generator never calls ``forward`` here — the whole stack runs inside the CUDA plan
(tts_b200/csrc).  ``forward`` is the autograd path training needs (out of scope).
"""
from torch import nn
from torch.nn.utils import weight_norm


def _wn_conv(channels, kernel_size, dilation=1):
    return weight_norm(nn.Conv1d(channels, channels, kernel_size=kernel_size, dilation=dilation, bias=True))


class ResidualStack(nn.Module):
    def __init__(self, channels, num_res_blocks, kernel_size):
        super().__init__()
        assert (kernel_size - 1) % 2 == 0, " [!] kernel_size has to be odd."
        half = (kernel_size - 1) // 2
        self.channels = channels
        self.kernel_size = kernel_size
        blocks, shortcuts = [], []
        # creation order (all blocks, then all shortcuts) fixes both the state-dict order and the
        # RNG stream of default init, so a seed reproduces the reference's weights
        for b in range(num_res_blocks):
            d = kernel_size ** b
            blocks.append(nn.Sequential(
                nn.LeakyReLU(0.2),
                nn.ReflectionPad1d(half * d),
                _wn_conv(channels, kernel_size, d),
                nn.LeakyReLU(0.2),
                _wn_conv(channels, 1),
            ))
        for _ in range(num_res_blocks):
            shortcuts.append(_wn_conv(channels, 1))
        self.blocks = nn.ModuleList(blocks)
        self.shortcuts = nn.ModuleList(shortcuts)

    def forward(self, x):
        for block, shortcut in zip(self.blocks, self.shortcuts):
            x = shortcut(x) + block(x)
        return x

    def remove_weight_norm(self):
        for m in self.modules():
            if isinstance(m, nn.Conv1d) and hasattr(m, "weight_g"):
                nn.utils.remove_weight_norm(m)
