"""Import shim for `diffusers` (absent from this image): the slice the example
scripts use — `DDPMScheduler` (DDPM, Ho et al. 2020: add_noise / set_timesteps /
step with `epsilon` or `sample` prediction, fixed_small variance), its DDIM
sibling, and the timestep embeddings.  Pipelines / UNet2DModel are names only."""
from types import SimpleNamespace

import numpy as np
import torch

from .models.embeddings import TimestepEmbedding, Timesteps  # noqa: F401

__version__ = "0.0-shim"


def _betas(num_train_timesteps, beta_start, beta_end, beta_schedule):
    if beta_schedule == "linear":
        return torch.linspace(beta_start, beta_end, num_train_timesteps, dtype=torch.float32)
    if beta_schedule == "scaled_linear":
        return torch.linspace(beta_start ** 0.5, beta_end ** 0.5, num_train_timesteps, dtype=torch.float32) ** 2
    if beta_schedule == "squaredcos_cap_v2":
        f = lambda t: np.cos((t + 0.008) / 1.008 * np.pi / 2) ** 2  # noqa: E731
        b = [min(1 - f((i + 1) / num_train_timesteps) / f(i / num_train_timesteps), 0.999)
             for i in range(num_train_timesteps)]
        return torch.tensor(b, dtype=torch.float32)
    raise NotImplementedError(beta_schedule)


class DDPMScheduler:
    def __init__(self, num_train_timesteps=1000, beta_start=0.0001, beta_end=0.02, beta_schedule="linear",
                 prediction_type="epsilon", clip_sample=True, variance_type="fixed_small", **kwargs):
        self.config = SimpleNamespace(num_train_timesteps=num_train_timesteps, beta_start=beta_start,
                                      beta_end=beta_end, beta_schedule=beta_schedule,
                                      prediction_type=prediction_type, clip_sample=clip_sample,
                                      variance_type=variance_type, **kwargs)
        self.betas = _betas(num_train_timesteps, beta_start, beta_end, beta_schedule)
        self.alphas = 1.0 - self.betas
        self.alphas_cumprod = torch.cumprod(self.alphas, dim=0)
        self.one = torch.tensor(1.0)
        self.init_noise_sigma = 1.0
        self.num_inference_steps = None
        self.timesteps = torch.arange(num_train_timesteps - 1, -1, -1)

    def set_timesteps(self, num_inference_steps, device=None):
        self.num_inference_steps = int(num_inference_steps)
        ratio = self.config.num_train_timesteps // self.num_inference_steps
        ts = (np.arange(0, self.num_inference_steps) * ratio).round()[::-1].copy().astype(np.int64)
        self.timesteps = torch.from_numpy(ts).to(device) if device is not None else torch.from_numpy(ts)

    def add_noise(self, original_samples, noise, timesteps):
        ac = self.alphas_cumprod.to(device=original_samples.device, dtype=original_samples.dtype)
        t = timesteps.to(original_samples.device)
        a = ac[t] ** 0.5
        s = (1 - ac[t]) ** 0.5
        while a.dim() < original_samples.dim():
            a, s = a.unsqueeze(-1), s.unsqueeze(-1)
        return a * original_samples + s * noise

    def _prev(self, t):
        n = self.num_inference_steps or self.config.num_train_timesteps
        return t - self.config.num_train_timesteps // n

    def step(self, model_output, timestep, sample, generator=None, return_dict=True):
        t = int(timestep)
        prev_t = self._prev(t)
        ac = self.alphas_cumprod
        a_t = ac[t].item()
        a_prev = ac[prev_t].item() if prev_t >= 0 else 1.0
        b_t, b_prev = 1 - a_t, 1 - a_prev
        cur_alpha = a_t / a_prev
        cur_beta = 1 - cur_alpha
        if self.config.prediction_type == "epsilon":
            x0 = (sample - b_t ** 0.5 * model_output) / a_t ** 0.5
        elif self.config.prediction_type == "sample":
            x0 = model_output
        elif self.config.prediction_type == "v_prediction":
            x0 = a_t ** 0.5 * sample - b_t ** 0.5 * model_output
        else:
            raise ValueError(self.config.prediction_type)
        if self.config.clip_sample:
            x0 = x0.clamp(-1, 1)
        c0 = a_prev ** 0.5 * cur_beta / b_t
        ct = cur_alpha ** 0.5 * b_prev / b_t
        prev = c0 * x0 + ct * sample
        if t > 0:
            var = max(b_prev / b_t * cur_beta, 1e-20)
            prev = prev + var ** 0.5 * torch.randn(model_output.shape, device=model_output.device,
                                                   dtype=model_output.dtype, generator=generator)
        return SimpleNamespace(prev_sample=prev, pred_original_sample=x0)


class DDIMScheduler(DDPMScheduler):
    def step(self, model_output, timestep, sample, eta=0.0, generator=None, return_dict=True):
        t = int(timestep)
        prev_t = self._prev(t)
        a_t = self.alphas_cumprod[t].item()
        a_prev = self.alphas_cumprod[prev_t].item() if prev_t >= 0 else 1.0
        if self.config.prediction_type == "epsilon":
            x0 = (sample - (1 - a_t) ** 0.5 * model_output) / a_t ** 0.5
            eps = model_output
        else:
            x0 = model_output
            eps = (sample - a_t ** 0.5 * x0) / (1 - a_t) ** 0.5
        prev = a_prev ** 0.5 * x0 + (1 - a_prev) ** 0.5 * eps
        return SimpleNamespace(prev_sample=prev, pred_original_sample=x0)


class _NameOnly:
    def __init__(self, *args, **kwargs):
        raise RuntimeError("diffusers shim: this class is a name only")


class DDPMPipeline(_NameOnly):
    pass


class UNet2DModel(_NameOnly):
    pass
