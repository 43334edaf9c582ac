"""Shrink AutoEncoder (SAE) and plain AutoEncoder as ``nn.Module`` front-ends.

These modules exist for API parity and host-side tooling (checkpoint I/O, the
CPU oracle, initialisation).  The hot path never runs them: training, scoring
and evaluation execute from the packed parameter slab inside the sm_100a kernels
(``csrc/``).  ``pack_model`` / ``load_packed`` convert between the two worlds.

Behavioural contract (reference: ``src/Model/Shrink_Autoencoder.py:21-167``,
``src/Model/AutoEncoder.py:21-159``):

* network ``D -> H -> ReLU -> Z`` / ``Z -> H -> ReLU -> D`` with biases, no output
  activation; ``state_dict`` keys ``encoder.encoder_network.{0,2}.*`` and
  ``decoder.decoder_network.{0,2}.*``;
* init ``W ~ U(-1/sqrt(fan_in), 1/sqrt(fan_in))``, ``b = 0`` and the same draw
  order on the torch generator as the reference (each ``nn.Linear`` draws its
  default init first, then the weights are re-drawn), so the same seed yields the
  same initial weights;
* ``forward(x) -> (latent, output, loss)``;
* SAE loss ``MSE_mean(x, x_hat) + lambda * mean_b ||z_b||_2`` with an *unsquared*
  norm (quirk Q15); AE loss is the MSE term only.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
from torch import nn

from ..layout import ParamLayout


def _reinit(seq: nn.Sequential) -> None:
    for mod in seq:
        if isinstance(mod, nn.Linear):
            lim = 1.0 / math.sqrt(mod.in_features)
            with torch.no_grad():
                mod.weight.uniform_(-lim, lim)
                mod.bias.zero_()


class Encoder(nn.Module):
    def __init__(self, input_dim: int = 40, hidden_neus: int = 27, latent_dim: int = 7):
        super().__init__()
        self.encoder_network = nn.Sequential(
            nn.Linear(input_dim, hidden_neus),
            nn.ReLU(),
            nn.Linear(hidden_neus, latent_dim),
        )
        self.init_params()

    def init_params(self) -> None:
        _reinit(self.encoder_network)

    def forward(self, inputs: torch.Tensor) -> torch.Tensor:
        return self.encoder_network(inputs)


class Decoder(nn.Module):
    def __init__(self, latent_dim: int = 7, hidden_neus: int = 27, output_dim: int = 40):
        super().__init__()
        self.decoder_network = nn.Sequential(
            nn.Linear(latent_dim, hidden_neus),
            nn.ReLU(),
            nn.Linear(hidden_neus, output_dim),
        )
        self.init_params()

    def init_params(self) -> None:
        _reinit(self.decoder_network)

    def forward(self, latent: torch.Tensor) -> torch.Tensor:
        return self.decoder_network(latent)


class _AEBase(nn.Module):
    model_type = "autoencoder"

    def __init__(self, input_dim: int, output_dim: int, hidden_neus: int, latent_dim: int):
        super().__init__()
        if input_dim != output_dim:
            raise ValueError("autoencoder requires input_dim == output_dim")
        self.encoder = Encoder(input_dim, hidden_neus, latent_dim)
        self.decoder = Decoder(latent_dim, hidden_neus, output_dim)
        self.layout = ParamLayout(input_dim, hidden_neus, latent_dim)

    # -- loss hooks --------------------------------------------------------
    def _loss(self, x, x_hat, z):  # pragma: no cover - abstract
        raise NotImplementedError

    def forward(self, input: torch.Tensor):
        latent = self.encoder(input)
        output = self.decoder(latent)
        return latent, output, self._loss(input, output, latent)

    # -- packed-slab bridge --------------------------------------------------
    def pack(self, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        return self.layout.pack(self.state_dict(), out)

    def load_packed(self, row: torch.Tensor) -> None:
        self.load_state_dict(self.layout.unpack(row))

    def _to_numpy(self, tensor: torch.Tensor):
        return tensor.detach().cpu().numpy()


class Shrink_Autoencoder(_AEBase):
    """SAE: reconstruction MSE plus ``shrink_lambda * mean ||z||_2``."""

    model_type = "hybrid"

    def __init__(self, input_dim: int = 40, output_dim: int = 40, hidden_neus: int = 27,
                 latent_dim: int = 7, shrink_lambda: float = 10):
        super().__init__(input_dim, output_dim, hidden_neus, latent_dim)
        self.shrink_lambda = shrink_lambda

    def shrink_loss(self, input, output, latent):
        recon = torch.mean((input - output) ** 2)
        shrink = torch.linalg.vector_norm(latent, dim=1).sum() / latent.shape[0]
        return recon + self.shrink_lambda * shrink

    _loss = shrink_loss


class Autoencoder(_AEBase):
    """Plain AE: reconstruction MSE only."""

    model_type = "autoencoder"

    def __init__(self, input_dim: int = 40, output_dim: int = 40, hidden_neus: int = 27,
                 latent_dim: int = 7):
        super().__init__(input_dim, output_dim, hidden_neus, latent_dim)

    def recon_loss(self, input, output, latent=None):
        return torch.mean((input - output) ** 2)

    _loss = recon_loss


def build_model(model_type: str, dim: int = 115, hidden: int = 27, latent: int = 7,
                shrink_lambda: float = 10) -> _AEBase:
    """Factory used by the engine: ``hybrid`` -> SAE, ``autoencoder`` -> AE."""
    if model_type == "hybrid":
        return Shrink_Autoencoder(dim, dim, hidden, latent, shrink_lambda)
    if model_type == "autoencoder":
        return Autoencoder(dim, dim, hidden, latent)
    raise ValueError(f"unknown model_type {model_type!r}")


def init_packed_models(n_clients: int, model_type: str, layout: ParamLayout,
                       shrink_lambda: float = 10) -> torch.Tensor:
    """Initialise ``n_clients`` models with the reference's RNG order.

    Returns a CPU ``[n_clients, layout.padded]`` fp32 slab.  One module is built
    per client in client order on the global torch generator, exactly as the
    reference builds its trainers (``src/main.py:225-236``).
    """
    slab = torch.zeros(n_clients, layout.padded, dtype=torch.float32)
    for c in range(n_clients):
        m = build_model(model_type, layout.dim, layout.hidden, layout.latent, shrink_lambda)
        m.pack(slab[c])
    return slab


def state_dict_from_packed(row: torch.Tensor, layout: ParamLayout) -> Dict[str, torch.Tensor]:
    return layout.unpack(row)
