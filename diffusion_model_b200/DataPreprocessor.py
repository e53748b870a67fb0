"""Drop-in for the reference's DataPreprocessor.py: ``SpectrumCompressor`` (DataPreprocessor.py:4-22), the per-node MLP
(Linear/ReLU stack) that turns a [N, spectrum_size] XANES spectrum into the compressed conditioning block of
h = [h_t | compressed spectrum | exO | t/T] (parts/train_per_iretation.py:136-145).

Same constructor, ``state_dict`` keys (``mlp.<i>.weight/bias``) and initialisation order as the reference, so
``model.pth['spectrum_compressor']`` checkpoints load unchanged.  The conditioning block does not depend on the diffusion
step: training evaluates it once per minibatch, the sampler once per trajectory (sampler.sample_structures keeps it
resident and the EGNN's first layer reads it with the rest of h).  Its three small dense layers are plain library GEMMs;
the EGNN that consumes the block runs on this package's kernels (fp32 mode for H = 36: the tcgen05 edge tiles hold
2H+1 <= 16 input slots).
"""
import torch
import torch.nn as nn


class SpectrumCompressor(nn.Module):
    def __init__(self, original_spectrum_dim, hidden_dim: list, compressed_spectrum_dim):
        super().__init__()
        assert isinstance(hidden_dim, list)
        self.original_spectrum_dim = original_spectrum_dim
        widths = [original_spectrum_dim] + list(hidden_dim)
        layers = []
        for a, b in zip(widths[:-1], widths[1:]):
            layers += [nn.Linear(a, b), nn.ReLU()]
        layers.append(nn.Linear(widths[-1], compressed_spectrum_dim))
        self.mlp = nn.Sequential(*layers)

    def forward(self, spectrum: torch.Tensor):
        assert spectrum.shape[1] == self.original_spectrum_dim
        return self.mlp(spectrum)
