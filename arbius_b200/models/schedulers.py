"""Diffusion schedulers as *linear update plans*.

Every sampler in the AnythingV3 template enum (reference: templates/anythingv3.json ``scheduler``:
DDIM, K_EULER, DPMSolverMultistep, K_EULER_ANCESTRAL, PNDM, KLMS) and the DDPM ancestral /
DDIM samplers Kandinsky and the video models use reduce, per step, to a few assignments

    dst = sum_k coef_k * src_k        src in {x, eps, noise, history buffers}

with coefficients known on the host in float64 before the loop starts. ``plan(i)`` returns those
assignments; the torch executor below is the CPU oracle and the native engine runs the same plan
with one fused kernel per assignment (``ab_lincomb``), so no sampler needs device-side control
flow and the whole denoising loop can be captured in a CUDA graph. Math follows the public
diffusers formulations for epsilon-prediction models.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch

Op = Tuple[str, List[Tuple[float, str]]]   # (dst, [(coef, src), ...])


def make_betas(n: int = 1000, beta_start: float = 0.00085, beta_end: float = 0.012,
               schedule: str = "scaled_linear") -> np.ndarray:
    if schedule == "scaled_linear":
        return np.linspace(beta_start ** 0.5, beta_end ** 0.5, n, dtype=np.float64) ** 2
    if schedule == "linear":
        return np.linspace(beta_start, beta_end, n, dtype=np.float64)
    if schedule == "cosine":   # improved-DDPM (Kandinsky prior)
        f = lambda t: math.cos((t + 0.008) / 1.008 * math.pi / 2) ** 2
        return np.array([min(1 - f((i + 1) / n) / f(i / n), 0.999) for i in range(n)], dtype=np.float64)
    raise ValueError(schedule)


class Scheduler:
    """Base: holds alphas_cumprod and the per-run timestep table."""
    name = "base"
    history: Sequence[str] = ()
    stochastic = False

    def __init__(self, num_train_timesteps: int = 1000, beta_start: float = 0.00085, beta_end: float = 0.012,
                 beta_schedule: str = "scaled_linear", steps_offset: int = 1):
        self.T = num_train_timesteps
        self.betas = make_betas(num_train_timesteps, beta_start, beta_end, beta_schedule)
        self.alphas_cumprod = np.cumprod(1.0 - self.betas)
        self.steps_offset = steps_offset
        self.timesteps: List[float] = []
        self.init_noise_sigma = 1.0

    # -- interface -------------------------------------------------------------------------
    def set_timesteps(self, n: int) -> None:
        raise NotImplementedError

    def input_scale(self, i: int) -> float:
        return 1.0

    def plan(self, i: int) -> List[Op]:
        raise NotImplementedError

    def __len__(self) -> int:
        return len(self.timesteps)

    # -- helpers ---------------------------------------------------------------------------
    def _k_sigmas(self, timesteps: np.ndarray) -> np.ndarray:
        sig = np.sqrt((1 - self.alphas_cumprod) / self.alphas_cumprod)
        return np.interp(timesteps, np.arange(self.T), sig)


class DDIM(Scheduler):
    name = "DDIM"

    def set_timesteps(self, n: int) -> None:
        ratio = self.T // n
        self.ratio = ratio
        self.timesteps = [float(t) for t in (np.arange(0, n) * ratio).round()[::-1].astype(np.int64) + self.steps_offset]

    def plan(self, i: int) -> List[Op]:
        t = int(self.timesteps[i])
        prev = t - self.ratio
        a_t = self.alphas_cumprod[t]
        a_p = self.alphas_cumprod[prev] if prev >= 0 else self.alphas_cumprod[0]
        cx = math.sqrt(a_p / a_t)
        ce = math.sqrt(1 - a_p) - math.sqrt(a_p) * math.sqrt(1 - a_t) / math.sqrt(a_t)
        return [("x", [(cx, "x"), (ce, "eps")])]


class PNDM(Scheduler):
    """PLMS (skip_prk_steps=True), as Stable Diffusion ships it."""
    name = "PNDM"
    history = ("e1", "e2", "e3", "xs", "ea")

    def set_timesteps(self, n: int) -> None:
        ratio = self.T // n
        self.ratio = ratio
        ts = (np.arange(0, n) * ratio).round().astype(np.int64) + self.steps_offset
        plms = np.concatenate([ts[:-1], ts[-2:-1], ts[-1:]])[::-1]
        self.timesteps = [float(t) for t in plms]

    def _prev(self, t: int, prev: int, sample: str, mo: List[Tuple[float, str]]) -> List[Tuple[float, str]]:
        a_t = self.alphas_cumprod[t]
        a_p = self.alphas_cumprod[prev] if prev >= 0 else self.alphas_cumprod[0]
        sc = math.sqrt(a_p / a_t)
        denom = a_t * math.sqrt(1 - a_p) + math.sqrt(a_t * (1 - a_t) * a_p)
        k = -(a_p - a_t) / denom
        return [(sc, sample)] + [(k * c, s) for c, s in mo]

    def plan(self, i: int) -> List[Op]:
        t = int(self.timesteps[i])
        prev = t - self.ratio
        ops: List[Op] = []
        # history after this call: e1 = newest eps. Counter semantics follow diffusers' step_plms.
        if i == 0:
            ops.append(("xs", [(1.0, "x")]))
            ops.append(("x", self._prev(t, prev, "x", [(1.0, "eps")])))
            ops.append(("e1", [(1.0, "eps")]))
        elif i == 1:
            # second call re-evaluates at the same nominal time: average with the first eps
            t2, prev2 = t + self.ratio, t
            ops.append(("ea", [(0.5, "eps"), (0.5, "e1")]))
            ops.append(("x", self._prev(t2, prev2, "xs", [(1.0, "ea")])))
        elif i == 2:
            ops.append(("x", self._prev(t, prev, "x", [(1.5, "eps"), (-0.5, "e1")])))
            ops += [("e2", [(1.0, "e1")]), ("e1", [(1.0, "eps")])]
        elif i == 3:
            ops.append(("x", self._prev(t, prev, "x", [(23 / 12, "eps"), (-16 / 12, "e1"), (5 / 12, "e2")])))
            ops += [("e3", [(1.0, "e2")]), ("e2", [(1.0, "e1")]), ("e1", [(1.0, "eps")])]
        else:
            ops.append(("x", self._prev(t, prev, "x", [(55 / 24, "eps"), (-59 / 24, "e1"), (37 / 24, "e2"),
                                                        (-9 / 24, "e3")])))
            ops += [("e3", [(1.0, "e2")]), ("e2", [(1.0, "e1")]), ("e1", [(1.0, "eps")])]
        return ops


class _KScheduler(Scheduler):
    def set_timesteps(self, n: int) -> None:
        ts = np.linspace(0, self.T - 1, n, dtype=np.float64)[::-1].copy()
        self.sigmas = np.concatenate([self._k_sigmas(ts), [0.0]])
        self.timesteps = [float(t) for t in ts]
        self.init_noise_sigma = float(self.sigmas.max())

    def input_scale(self, i: int) -> float:
        return 1.0 / math.sqrt(self.sigmas[i] ** 2 + 1.0)


class Euler(_KScheduler):
    name = "K_EULER"

    def plan(self, i: int) -> List[Op]:
        return [("x", [(1.0, "x"), (self.sigmas[i + 1] - self.sigmas[i], "eps")])]


class EulerAncestral(_KScheduler):
    name = "K_EULER_ANCESTRAL"
    stochastic = True

    def plan(self, i: int) -> List[Op]:
        s_from, s_to = self.sigmas[i], self.sigmas[i + 1]
        s_up = math.sqrt(max(s_to ** 2 * (s_from ** 2 - s_to ** 2) / s_from ** 2, 0.0))
        s_down = math.sqrt(max(s_to ** 2 - s_up ** 2, 0.0))
        return [("x", [(1.0, "x"), (s_down - s_from, "eps"), (s_up, "noise")])]


class LMS(_KScheduler):
    name = "KLMS"
    history = ("d1", "d2", "d3")
    order = 4

    def _coeff(self, order: int, i: int, k: int) -> float:
        """Integral over [sigma_i, sigma_{i+1}] of the k-th Lagrange basis polynomial (exact)."""
        nodes = [self.sigmas[i - j] for j in range(order)]
        poly = np.poly1d([1.0])
        for j in range(order):
            if j == k:
                continue
            poly = poly * np.poly1d([1.0, -nodes[j]]) / (nodes[k] - nodes[j])
        integ = poly.integ()
        return float(integ(self.sigmas[i + 1]) - integ(self.sigmas[i]))

    def plan(self, i: int) -> List[Op]:
        order = min(i + 1, self.order)
        names = ["eps", "d1", "d2", "d3"][:order]
        terms = [(1.0, "x")] + [(self._coeff(order, i, k), names[k]) for k in range(order)]
        ops: List[Op] = [("x", terms)]
        if i >= 2:
            ops.append(("d3", [(1.0, "d2")]))
        if i >= 1:
            ops.append(("d2", [(1.0, "d1")]))
        ops.append(("d1", [(1.0, "eps")]))
        return ops


class DPMSolverMultistep(Scheduler):
    """DPM-Solver++(2M), midpoint, final sigma 0 => first-order last step."""
    name = "DPMSolverMultistep"
    history = ("m1",)

    def set_timesteps(self, n: int) -> None:
        ts = np.linspace(0, self.T - 1, n + 1).round()[::-1][:-1].astype(np.int64)
        self.sigmas = np.concatenate([self._k_sigmas(ts.astype(np.float64)), [0.0]])
        self.timesteps = [float(t) for t in ts]

    @staticmethod
    def _alpha_sigma(sig: float) -> Tuple[float, float]:
        a = 1.0 / math.sqrt(sig * sig + 1.0)
        return a, sig * a

    def plan(self, i: int) -> List[Op]:
        n = len(self.timesteps)
        a_s, s_s = self._alpha_sigma(self.sigmas[i])
        a_t, s_t = self._alpha_sigma(self.sigmas[i + 1])
        lam_s = math.log(a_s) - math.log(s_s)
        last = i == n - 1
        # x0 prediction of this step: m0 = x/a_s - (s_s/a_s) eps
        m0 = [(1.0 / a_s, "x"), (-s_s / a_s, "eps")]
        if last:   # sigma_t == 0: x <- x0
            return [("x", m0)]
        lam_t = math.log(a_t) - math.log(s_t)
        h = lam_t - lam_s
        k = -a_t * (math.exp(-h) - 1.0)
        ops: List[Op] = []
        if i == 0:
            ops.append(("m1", m0))
            ops.append(("x", [(s_t / s_s, "x"), (k, "m1")]))
            return ops
        a_p, s_p = self._alpha_sigma(self.sigmas[i - 1])
        lam_p = math.log(a_p) - math.log(s_p)
        r0 = (lam_s - lam_p) / h
        # D = (1 + 1/(2 r0)) m0 - 1/(2 r0) m1
        w0, w1 = 1.0 + 0.5 / r0, -0.5 / r0
        terms = [(s_t / s_s + k * w0 * m0[0][0], "x"), (k * w0 * m0[1][0], "eps"), (k * w1, "m1")]
        ops.append(("tmp", terms))
        ops.append(("m1", m0))
        ops.append(("x", [(1.0, "tmp")]))
        return ops


class DDPMAncestral(Scheduler):
    """DDPM ancestral sampling with fixed-small variance on a strided timestep grid
    (Kandinsky ``p_sampler``; fresh noise every step)."""
    name = "p_sampler"
    stochastic = True

    def set_timesteps(self, n: int) -> None:
        ts = np.linspace(0, self.T - 1, n).round()[::-1].astype(np.int64)
        self.timesteps = [float(t) for t in ts]

    def plan(self, i: int) -> List[Op]:
        t = int(self.timesteps[i])
        prev = int(self.timesteps[i + 1]) if i + 1 < len(self.timesteps) else -1
        a_t = self.alphas_cumprod[t]
        a_p = self.alphas_cumprod[prev] if prev >= 0 else 1.0
        beta = 1.0 - a_t / a_p
        # posterior mean in terms of (x, eps)
        c_x0 = math.sqrt(a_p) * beta / (1 - a_t)
        c_xt = math.sqrt(a_t / a_p) * (1 - a_p) / (1 - a_t)
        cx = c_xt + c_x0 / math.sqrt(a_t)
        ce = -c_x0 * math.sqrt(1 - a_t) / math.sqrt(a_t)
        var = beta * (1 - a_p) / (1 - a_t) if prev >= 0 else 0.0
        terms = [(cx, "x"), (ce, "eps")]
        if var > 0:
            terms.append((math.sqrt(var), "noise"))
        return [("x", terms)]


SCHEDULERS = {c.name: c for c in (DDIM, PNDM, Euler, EulerAncestral, LMS, DPMSolverMultistep, DDPMAncestral)}
SCHEDULERS.update({"ddim_sampler": DDIM, "p_sampler": DDPMAncestral, "plms_sampler": PNDM, "pims_sampler": PNDM})


def make_scheduler(name: str, **kw) -> Scheduler:
    if name not in SCHEDULERS:
        raise ValueError(f"unknown scheduler {name!r}; choices: {sorted(SCHEDULERS)}")
    return SCHEDULERS[name](**kw)


def apply_plan_torch(ops: List[Op], bufs: Dict[str, torch.Tensor]) -> None:
    """Reference executor (fp32 torch tensors; left-to-right accumulation like the kernel)."""
    for dst, terms in ops:
        acc = None
        for c, s in terms:
            v = bufs[s] * float(np.float32(c))
            acc = v if acc is None else acc + v
        bufs[dst] = acc
