#!/usr/bin/env python
"""bench.py -- the VMC hot path of the 2D Heisenberg 10x10 tensorized-GRU wavefunction (BASELINE.json
config 4, the configuration the north-star target is quoted on), one process per GPU.

A "step" is one full VMC step over one batch of synthetic input (random-init weights):
    sample 10^4 configurations  ->  enumerate + evaluate the ~1.3*10^6 connected configurations and reduce
    E_loc  ->  energy moments (all-reduce)  ->  backward of log psi seeded with (E_loc - <E>)
    ->  gradient all-reduce  ->  TF-Adam update  ->  weight repack.
metric  = E_loc configurations per second (1 diagonal + every connected configuration with non-zero matrix
          element, summed over samples) through that whole step, whole job.   samples/s is reported beside it.
value   : inputs resident in HBM (Philox uniforms generated in-kernel).
e2e     : same step through the reference-facing API with HOST buffers: pinned uniforms H2D, samples and
          local energies D2H (as the reference's sess.run fetches), Eloc fed back H2D for the backward.
roofline: the dominant kernel (connected-configuration sweep).  The headline runs the tcgen05 tensor-core regime with
          FP16 operands (north_star's reduced-precision mode, 1e-3 bound): tensor bound, executed MMA flops over the
          measured sustained 16-bit peak of MEASURED_PEAKS.json.  The split-precision regime (tc3, 1e-5 bound) and the
          FFMA regime (fp32, bound = FFMA peak measured live by a probe kernel) are reported under "workloads" together
          with the other configurations of BASELINE.json (N = 1).
scaling : "weak" (numsamples per GPU fixed) by default; at N > 1 the line also carries "strong_scaling" (the N = 1
          batch sharded over the ranks); --scaling strong makes that the headline instead.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME] [--precision P] [--scaling S]
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

WORKLOADS = {
    # name: (engine kwargs, hamiltonian name, ham params builder, numsamples, description)
    "heis10x10_tri_tgru100": dict(eng=dict(cell="TENSGRU2D", nx=10, ny=10, num_units=100, complex_head=True, u1=True),
                                  ham="HEIS2D", hp=[1.0], ns=10000,
                                  desc="2DHeisenberg 10x10 triangular (J2=1) tensorized complex GRU, 100 units, U(1) Sz=0, 1e4 samples/step/GPU"),
    "heis10x10_sq_tgru100": dict(eng=dict(cell="TENSGRU2D", nx=10, ny=10, num_units=100, complex_head=True, u1=True),
                                 ham="HEIS2D", hp=[0.0], ns=10000,
                                 desc="2DHeisenberg 10x10 square (J2=0) tensorized complex GRU, 100 units, U(1) Sz=0, 1e4 samples/step/GPU"),
    "tfim1d_n20_gru50": dict(eng=dict(cell="GRU1D", nx=20, ny=1, num_units=50, num_layers=1), ham="TFIM1D",
                             hp=[1.0] + [1.0] * 19, ns=500, desc="1DTFIM N=20 pRNN GRU 1x50, 500 samples/step/GPU"),
    "tfim1d_n20_gru50x3": dict(eng=dict(cell="GRU1D", nx=20, ny=1, num_units=50, num_layers=3), ham="TFIM1D",
                               hp=[1.0] + [1.0] * 19, ns=500, desc="1DTFIM N=20 pRNN GRU 3x50 (BASELINE.json config 1), 500 samples/step/GPU"),
    "j1j2_n20_cgru50": dict(eng=dict(cell="GRU1D", nx=20, ny=1, num_units=50, complex_head=True, u1=True), ham="J1J2",
                            hp=[0.0, 0.0] + [1.0] * 20 + [0.2] * 20 + [0.0] * 20, ns=500,
                            desc="J1J2 N=20 cRNN GRU 50, J2=0.2, U(1), 500 samples/step/GPU"),
    "tfim2d_12x12_rnn50": dict(eng=dict(cell="VANILLA2D", nx=12, ny=12, num_units=50), ham="TFIM2D",
                               hp=[3.0] + [1.0] * 144, ns=10000, desc="2DTFIM 12x12 2D RNN 50 units, h=3, 1e4 samples/step/GPU"),
    "rydberg_4x4_gru256": dict(eng=dict(cell="GRU4", nx=4, ny=4, num_units=256), ham="RYDBERG",
                               hp=[1.7 ** 6, 3.3, 1.0, 2.001], ns=12500,
                               desc="Rydberg Kagome 4x4 cells (48 atoms) 2D GRU 256 units, 12500 samples/step/GPU"),
    "rydberg_8x8_gru256": dict(eng=dict(cell="GRU4", nx=8, ny=8, num_units=256), ham="RYDBERG",
                               hp=[1.7 ** 6, 3.3, 1.0, 2.001], ns=12500,
                               desc="Rydberg Kagome 8x8 cells (192 atoms) 2D GRU 256 units, 12500 samples/step/GPU"),
}
DEFAULT_WORKLOAD = "heis10x10_tri_tgru100"


# ------------------------------------------------------------------------------------------------------
# algorithmic / executed work
# ------------------------------------------------------------------------------------------------------
def flops_per_site(w):
    """Algorithmic flops per configuration per site (2 flop per multiply-add, activations excluded)."""
    e = w["eng"]
    nh = e["num_units"]
    cell = e["cell"]
    if cell == "TENSGRU2D":
        return 2 * (2 * nh * 4) + 2 * 2 * (8 * nh * nh) + 2 * (2 * nh * nh) + 2 * (2 * nh * 2)   # 362 400 at nh=100
    if cell == "GRU1D":
        heads = 2 if e.get("complex_head") else 1
        L = e.get("num_layers", 1)
        return 6 * nh * (2 + nh) + (L - 1) * 6 * nh * (nh + nh) + heads * 2 * nh * 2
    if cell == "VANILLA2D":
        return 2 * (2 * nh * nh + 4 * nh) + 4 * nh
    if cell == "GRU4":
        K = 4 * (nh + 6)
        return 2 * 2 * K * nh + 2 * 4 * nh * nh + 2 * nh * 8
    raise ValueError(cell)


def executed_flops_per_step_site(w):
    """Flops the sweep kernel actually executes per row per interior site (one-hot inputs are row selects,
    tensorized weight slices are pre-summed per class): 2*K*N_out + heads."""
    e = w["eng"]
    nh = e["num_units"]
    cell = e["cell"]
    if cell == "TENSGRU2D":
        return 2 * (2 * nh) * (3 * nh) + 2 * nh * 4
    if cell == "GRU1D":
        heads = 2 if e.get("complex_head") else 1
        L = e.get("num_layers", 1)
        return 2 * nh * 3 * nh + (L - 1) * 2 * (2 * nh) * 3 * nh + heads * 2 * nh * 2
    if cell == "VANILLA2D":
        return 2 * (2 * nh) * nh + 2 * nh * 2
    if cell == "GRU4":
        return 2 * (2 * nh) * (3 * nh) + 2 * nh * 8     # two of the four neighbour slots are live on interior sites
    raise ValueError(cell)


# ------------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons of one GPU, sampled every 25 ms during the timed region.

    NVML in a background thread (per-device queries, microseconds each).  An `nvidia-smi -lms` child process is only the
    fallback: on an 8-GPU box it enumerates every GPU per poll and serialises on the driver, which stalled this process's own
    kernel launches (C4 backward phase 7 -> 24 ms with one poller, 33 -> 43 ms/step with one poller per rank)."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.p = None
        self.f = None
        self.thread = None
        self.samples = []          # (sm_mhz, max_mhz, reasons bitmask)
        self.nvml = None
        self.handle = None
        try:
            import pynvml
            pynvml.nvmlInit()
            pr = torch.cuda.get_device_properties(gpu_index)
            try:
                bus = "%08X:%02X:%02X.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
                self.handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
            except Exception:
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                phys = int(vis.split(",")[gpu_index]) if vis and vis.split(",")[gpu_index].strip().isdigit() else gpu_index
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nvml = pynvml
            # the first queries initialise NVML lazily (hundreds of ms, under a driver lock that kernel launches also take):
            # make them here, before the warm-up steps, not inside the timed region
            pynvml.nvmlDeviceGetClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
            pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
            try:
                pynvml.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
            except Exception:
                pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
        except Exception:
            self.nvml = None

    def _poll(self):
        nv = self.nvml
        while not self._stop:
            try:
                sm = nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)
                mx = nv.nvmlDeviceGetMaxClockInfo(self.handle, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                self.samples.append((float(sm), float(mx), int(rs)))
            except Exception:
                pass
            time.sleep(0.025)

    def start(self):
        if self.nvml is not None:
            import threading
            self._stop = False
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        try:
            self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.FIELDS}",
                                       "--format=csv,noheader,nounits", "-lms", "25"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.thread is not None:
            self._stop = True
            self.thread.join(timeout=2)
            nv = self.nvml
            bits = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}
            reasons = set()
            for _, _, rs in self.samples:
                for name, bit in bits.items():
                    if rs & bit:
                        reasons.add(name)
            if self.samples:
                out.update(sm_mhz=float(np.median([x[0] for x in self.samples])), sm_max_mhz=float(max(x[1] for x in self.samples)),
                           reasons=sorted(reasons), samples=len(self.samples), source="nvml")
            return out
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm), source="nvidia-smi")
        return out


# ------------------------------------------------------------------------------------------------------
# CPU baseline: the float64 oracle (kind "port"; TensorFlow 1.13 is unavailable on this box)
# ------------------------------------------------------------------------------------------------------
def oracle_step(w, ns, seed=0):
    """One VMC step of the workload on the host cores with the oracle: sample + E_loc + gradient.
    Returns (configs evaluated, seconds, breakdown)."""
    from oracle import wavefunctions as owf, local_energies as ole, costs
    e = w["eng"]
    cell = e["cell"]
    nh, nx, ny = e["num_units"], e["nx"], e["ny"]
    g = torch.Generator().manual_seed(seed)
    P = make_oracle_params(w, g)
    nsteps = nx * ny
    u = torch.rand(ns, nsteps, generator=g, dtype=torch.float32)
    cname = {"VANILLA2D": "vanilla", "TENS2D": "tens", "TENSGRU2D": "tensgru"}

    def run(samples=None, grad=False):
        ctx = torch.enable_grad() if grad else torch.no_grad()
        with ctx:
            if cell == "GRU1D":
                return owf.rnn1d(P, nx, [nh] * e.get("num_layers", 1), samples=samples, uniforms=u, numsamples=ns,
                                 complex_=bool(e.get("complex_head")), u1=bool(e.get("u1")))
            if cell == "GRU4":
                return owf.rydberg2d(P, nx, ny, nh, samples=samples, uniforms=u, numsamples=ns)
            return owf.rnn2d(P, nx, ny, nh, cname[cell], samples=samples, uniforms=u, numsamples=ns,
                             complex_=bool(e.get("complex_head")), mag_fixed=bool(e.get("u1")))

    t0 = time.perf_counter()
    s = run()["samples"]
    t1 = time.perf_counter()

    def logpsi(cfg):
        r = run(samples=torch.as_tensor(cfg))
        return r["logp"].numpy(), r["phase"].numpy()

    sn = s.numpy()
    hp = w["hp"]
    ham = w["ham"]
    if ham == "HEIS2D":
        E, counts = ole.heis2d_local_energies(hp[0], sn, logpsi)
        nconf = int(counts.sum())
        if hp[0] == 0.0:   # count only the non-zero matrix elements, like the GPU metric
            nconf = ns + sum(int((ole.heis2d_connected(0.0, x)[1][1:] != 0).sum()) for x in sn)
    elif ham == "J1J2":
        N = nx
        E, counts = ole.j1j2_local_energies(np.asarray(hp[2:2 + N]), np.asarray(hp[2 + N:2 + 2 * N]), np.asarray(hp[2 + 2 * N:2 + 3 * N]),
                                            sn, logpsi, periodic=bool(hp[0]), Marshall_sign=bool(hp[1]))
        nconf = int(counts.sum())
    elif ham == "TFIM1D":
        E = ole.ising1d_local_energies(np.asarray(hp[1:] + [0.0]), hp[0], sn, logpsi)
        nconf = ns * (nx + 1)
    elif ham == "TFIM2D":
        E = ole.ising2d_local_energies(np.asarray(hp[1:]).reshape(nx, ny), hp[0], sn, logpsi)
        nconf = ns * (nx * ny + 1)
    else:
        pos = ole.generate_kagome(nx, ny)
        nb, dist = ole.neighbouring_atoms(pos, hp[3], nx, ny)
        E = ole.rydberg_local_energies(hp[0], hp[1], hp[2], nb, dist, sn, logpsi)
        nconf = ns * (3 * nx * ny + 1)
    t2 = time.perf_counter()
    for v in P.values():
        v.requires_grad_(True)
    r = run(samples=s, grad=True)
    Et = torch.as_tensor(E)
    if e.get("complex_head"):
        cost = costs.cost_complex(r["logp"], r["phase"], Et.to(torch.complex128))
    else:
        cost = costs.cost_real(r["logp"], Et.real if Et.is_complex() else Et)
    cost.backward()
    t3 = time.perf_counter()
    return nconf, t3 - t0, dict(sample_s=t1 - t0, eloc_s=t2 - t1, grad_s=t3 - t2)


def make_oracle_params(w, g):
    """Random-init parameters with the shapes of the TF variables (values do not matter for timing)."""
    e = w["eng"]
    nh, cell = e["num_units"], e["cell"]
    D = 8 if cell == "GRU4" else 2
    r = lambda *s: (torch.rand(*s, generator=g, dtype=torch.float64) - 0.5) * 0.4
    P = {"dense.W": r(nh, D), "dense.b": r(D)}
    if e.get("complex_head"):
        P["dense_phase.W"], P["dense_phase.b"] = r(nh, 2), r(2)
    if cell == "GRU1D":
        for l in range(e.get("num_layers", 1)):
            din = 2 if l == 0 else nh
            P.update({f"gru{l}.Wg": r(din + nh, 2 * nh), f"gru{l}.bg": r(2 * nh), f"gru{l}.Wci": r(din, nh),
                      f"gru{l}.Wch": r(nh, nh), f"gru{l}.bci": r(nh), f"gru{l}.bch": r(nh)})
    elif cell == "VANILLA2D":
        P.update(Wh=r(nh, nh), Uh=r(2, nh), Wv=r(nh, nh), Uv=r(2, nh), b=r(nh))
    elif cell in ("TENS2D", "TENSGRU2D"):
        P.update(W=r(nh, 2 * nh, 4), b=r(nh), Wg=r(nh, 2 * nh, 4), bg=r(nh), Wmerge=r(2 * nh, nh))
    else:
        P.update(W=r(4 * (nh + 6), nh), b=r(nh), Wg=r(4 * (nh + 6), nh), bg=r(nh), Wmerge=r(4 * nh, nh))
    return P


def cpu_baseline(w, target_s=15.0):
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n0 = 8
    nconf, dt, _ = oracle_step(w, n0)          # also warms the BLAS threads up
    nconf, dt, _ = oracle_step(w, n0)
    ns = int(max(n0, min(8192, n0 * target_s / max(dt, 1e-3))))
    nconf, dt, br = oracle_step(w, ns)
    return dict(value=nconf / dt, unit="eloc_configs/s", cores=cores, kind="port",
                sample=f"oracle (torch float64, {cores} threads) on {ns} samples = {nconf} configurations of the same workload: "
                       f"sample {br['sample_s']:.2f}s + E_loc (host enumeration + chunked forwards) {br['eloc_s']:.2f}s + autograd {br['grad_s']:.2f}s",
                seconds=dt)


# ------------------------------------------------------------------------------------------------------
# measurement of one workload on this rank's GPU
# ------------------------------------------------------------------------------------------------------
DTYPE = {"tc": "f16 operands / f32 accumulate (tcgen05); heads, log psi, E_loc in f32/f64",
         "tc3": "f16 hi+lo operand pairs (22 mantissa bits), 3 MMAs per K step, f32 accumulate (tcgen05); heads, log psi, E_loc in f32/f64",
         "fp32": "f32"}
# default regime per cell: the fast tensor-core regime (1e-3 bound) where it holds that bound at the benchmarked size
# (tests/test_gpu_bench_size.py), the split-precision one for the vanilla 2D RNN (plain FP16 operands measure ~1e-2 at
# 12x12, tools/cmp_regimes.py), FFMA for the 1D models
AUTO_PRECISION = {"TENSGRU2D": "tc", "TENS2D": "tc", "GRU4": "tc", "VANILLA2D": "tc3", "GRU1D": "fp32"}


def ev():
    return torch.cuda.Event(enable_timing=True)


class Bench:
    """One workload in one precision regime on the current device: the resident step, its phases and its roofline."""

    def __init__(self, name, precision, ns=None, prefix_reuse=True):
        from rnnwavefunctions_b200 import _lib, vmc
        from rnnwavefunctions_b200.engine import Engine
        self._lib = _lib
        self.name = name
        self.w = dict(WORKLOADS[name])
        if ns:
            self.w["ns"] = int(ns)
        self.ns = self.w["ns"]
        ek = dict(self.w["eng"])
        ek["cell"] = getattr(_lib, "CELL_" + ek["cell"])
        self.precision = AUTO_PRECISION[self.w["eng"]["cell"]] if precision == "auto" else precision
        self.eng = Engine(seed=111, precision=self.precision, **ek)
        self.ham = getattr(_lib, "HAM_" + self.w["ham"])
        self.V = vmc.VMC(self.eng, self.ham, self.w["hp"], self.ns, lr=5e-4, seed=111, prefix_reuse=prefix_reuse)

    def step(self):
        eng, V, ns = self.eng, self.V, self.ns
        e0, e1, e2, e3 = ev(), ev(), ev(), ev()
        offset = (V.it * V.world + V.rank) * ns
        e0.record()
        s, lp, ph = eng.sample(ns, seed=111, row_offset=offset, saved=V.saved)
        e1.record()
        e, cnt = eng.local_energies(self.ham, self.w["hp"], s, lp, ph, saved=V.saved if V.prefix_reuse else None, return_count=True)
        e2.record()
        grad, mean_e, var = V.gradient(s, lp, ph, e)
        V.opt.step(eng.params, grad, V.lr)
        eng.set_weights()
        e3.record()
        V.it += 1
        return (e0, e1, e2, e3), cnt, s

    def run(self, steps, warmup, barrier, clocks=None):
        """W untimed + K timed steps, CUDA events on the launching stream, barrier + synchronize on both sides."""
        for _ in range(warmup):
            self.step()
        barrier()
        launches0 = self._lib.lib.rnnwf_launch_count()
        if clocks is not None:
            clocks.start()
        t_start, t_end = ev(), ev()
        recs = []
        barrier()
        t_start.record()
        for _ in range(steps):
            recs.append(self.step())
        t_end.record()
        barrier()
        clk = clocks.stop() if clocks is not None else None
        out = {"elapsed_ms": t_start.elapsed_time(t_end), "launches": int(self._lib.lib.rnnwf_launch_count() - launches0),
               "phase_ms": {"sample": 0.0, "eloc": 0.0, "backward": 0.0}, "nconf": 0, "clocks": clk, "steps": steps}
        for (e0, e1, e2, e3), cnt, s in recs:
            out["phase_ms"]["sample"] += e0.elapsed_time(e1) / steps
            out["phase_ms"]["eloc"] += e1.elapsed_time(e2) / steps
            out["phase_ms"]["backward"] += e2.elapsed_time(e3) / steps
            out["nconf"] += int(cnt) + self.ns
        out["s_last"], out["nchild"] = recs[-1][2], int(recs[-1][1])
        return out

    # ---- roofline of the dominant kernel (the connected-configuration sweep inside local_energies)
    def roofline(self, res, peaks, ffma_peak):
        w, eng, V, precision, ns = self.w, self.eng, self.V, self.precision, self.ns
        eloc_ms = res["phase_ms"]["eloc"]
        nchild = res["nchild"]
        T = eng.nsteps
        exec_sites = executed_sites(w, res["s_last"].cpu().numpy(), eng) if V.prefix_reuse else nchild * T
        exec_flops = exec_sites * executed_flops_per_step_site(w)
        alg_flops = nchild * T * flops_per_site(w)
        bf16_peak = peaks.get("bf16_tflops_sustained", 1400.0)
        peak_src = ("MEASURED_PEAKS.json bf16_tflops_sustained (16-bit tensor peak, kernel timed inside a long step)"
                    if peaks else "fallback 1400 (B200_PROFILING.md)")
        common = {"algorithmic_flops_per_launch": alg_flops, "algorithmic_tflops": alg_flops / (eloc_ms * 1e-3) / 1e12,
                  "ffma_peak_tflops_measured": ffma_peak, "launch_ms": eloc_ms}
        if precision in ("tc", "tc3") and w["eng"]["cell"] == "GRU4":
            # sweep_tc4_kernel: per row and unit cell the dense [(live neighbours)*nhq + 32] x [3*nhq] contraction, nhq = 64*passes;
            # a child flipped in cell t re-runs cells t..T-1 (prefix reuse), three children per cell and sample
            nx, ny, nh = w["eng"]["nx"], w["eng"]["ny"], w["eng"]["num_units"]
            nhq = (nh + 63) // 64 * 64
            stepi = lambda ix, iy: iy * nx + (ix if iy % 2 == 0 else nx - 1 - ix)
            site_flops = []
            for iy in range(ny):
                for k in range(nx):
                    ix = k if iy % 2 == 0 else nx - 1 - k
                    d = 1 if iy % 2 == 0 else -1
                    t = iy * nx + k
                    keys = [(ix - d, iy), (ix, iy - 1), ((ix + d) % nx, iy), (ix, (iy + 1) % ny)]
                    live = sum(1 for (a, b) in keys if 0 <= a < nx and 0 <= b < ny and stepi(a, b) < t)
                    # split precision: hi hi + lo hi + hi lo on the hidden jobs, hi + lo weights on the exact one-hot input job
                    mh, mx = (3, 2) if precision == "tc3" else (1, 1)
                    site_flops.append(2 * (mh * live * nhq + mx * 32) * 3 * nhq + 2 * nh * 8)
            suffix = np.cumsum(site_flops[::-1])[::-1]
            per_sample = 3 * float(suffix.sum()) if V.prefix_reuse else 3 * T * float(suffix[0])
            exec_flops = ns * per_sample
            achieved = exec_flops / (eloc_ms * 1e-3) / 1e12
            return dict(common, bound="tensor", kernel=f"sweep_tc4_kernel (MODE_CHILD{', split precision' if precision == 'tc3' else ''})", achieved=achieved, peak=bf16_peak, unit="TFLOP/s",
                        frac=achieved / bf16_peak, traffic=None, peak_source=peak_src, executed_flops_per_launch=exec_flops,
                        mma_columns_per_row_site=3 * nhq,
                        note="executed = MMA flops issued per row (absent neighbours skipped, hidden axis padded to 64 per pass, K=32 "
                             "input/bias job, prefix reuse counted per row; split precision: three MMAs per K step, all counted); launch_ms brackets rnnwf_local_energies (enumeration kernels + sweep)")
        if precision in ("tc", "tc3"):
            # the tensor-core sweep executes, per row and site, the dense [2*nhp] x [7*16*passes] contraction
            nhp = (w["eng"]["num_units"] + 7) // 8 * 8
            npass = (w["eng"]["num_units"] + 15) // 16
            subs_last = (w["eng"]["num_units"] - 16 * (npass - 1) + 3) // 4       # the last column pass is issued only as wide as its real units
            ncols = 112 * (npass - 1) + (28 * subs_last + 15) // 16 * 16
            kname = "sweep_tc_kernel"
            if w["eng"]["cell"] == "VANILLA2D":
                ncols = 112                                                       # one pass per site: 7 unit pieces x 4 threads x 4 units
                if "two-tile sweep" in eng.describe():                            # sweep_van2_kernel: units in natural order, N = 64
                    ncols, kname = 64, "sweep_van2_kernel"
            mma_per_kstep = 3 if precision == "tc3" else 1          # split precision: A_hi B_hi + A_lo B_hi + A_hi B_lo
            exec_flops = exec_sites * (mma_per_kstep * 2 * (2 * nhp) * ncols + 2 * w["eng"]["num_units"] * 4)
            achieved = exec_flops / (eloc_ms * 1e-3) / 1e12
            return dict(common, bound="tensor", kernel=f"{kname} (MODE_CHILD{', split precision' if precision == 'tc3' else ''})",
                        achieved=achieved, peak=bf16_peak, unit="TFLOP/s", frac=achieved / bf16_peak, traffic=None, peak_source=peak_src,
                        executed_flops_per_launch=exec_flops, mma_columns_per_row_site=ncols,
                        note="executed = MMA flops actually issued (linearised class trick: 7 accumulator blocks per unit, narrow last pass; "
                             "split precision: three MMAs per K step, all counted; prefix reuse counted); launch_ms brackets "
                             "rnnwf_local_energies (enumeration kernels + sweep; the sweep is >99%)")
        achieved = exec_flops / (eloc_ms * 1e-3) / 1e12
        return dict(common, bound="fp32-ffma", kernel="sweep_kernel (MODE_CHILD)", achieved=achieved, peak=ffma_peak, unit="TFLOP/s",
                    frac=achieved / ffma_peak if ffma_peak > 0 else None, traffic=None,
                    peak_source="measured live: rnnwf_ffma_probe (16 FMA chains/thread, 592x512 threads), best of 5",
                    executed_flops_per_launch=exec_flops, frac_of_bf16_tensor_peak=achieved / bf16_peak,
                    note="launch_ms brackets rnnwf_local_energies (enumeration kernels + sweep; the sweep is >99% of it)")

    # ---- end to end through the reference-facing API with host buffers
    def e2e(self, K, barrier, world, dev):
        import torch.distributed as dist
        eng, V, w, ns, ham = self.eng, self.V, self.w, self.ns, self.ham
        pinned_u = torch.rand(ns, eng.nsteps, dtype=torch.float32).pin_memory()
        host_s = torch.empty((ns,) + eng.cfg_shape, dtype=torch.int8).pin_memory()
        host_e = torch.empty(ns, dtype=torch.complex128).pin_memory()
        host_eloc_in = torch.empty(ns, dtype=torch.complex128).pin_memory()
        h2d = pinned_u.numel() * 4 + host_eloc_in.numel() * 16 + host_s.numel()
        d2h = host_s.numel() + host_e.numel() * 16 + 16
        nconf = 0
        a, b = ev(), ev()
        for it in range(K + 1):                            # iteration 0: untimed warm-up of this path (pinned buffers, allocator)
            if it == 1:
                nconf = 0
                barrier()
                a.record()
            u = pinned_u.to(dev, non_blocking=True)
            s, lp, ph = eng.sample(ns, uniforms=u, saved=V.saved)
            e, cnt = eng.local_energies(ham, w["hp"], s, lp, ph, saved=V.saved if V.prefix_reuse else None, return_count=True)
            host_s.copy_(s, non_blocking=True)
            host_e.copy_(e, non_blocking=True)
            torch.cuda.synchronize()                       # the reference's sess.run returns numpy arrays here
            host_eloc_in.copy_(host_e)                     # meanE / varE on the host, Eloc fed back (optstep feed_dict)
            _ = complex(host_e.mean())
            s_dev = host_s.to(dev, non_blocking=True)
            e_dev = host_eloc_in.to(dev, non_blocking=True)
            grad, mean_e, var = V.gradient(s_dev, lp, ph, e_dev)
            V.opt.step(eng.params, grad, V.lr)
            eng.set_weights()
            _ = complex(mean_e)                            # D2H of the step's result
            nconf += int(cnt) + ns
        b.record()
        barrier()
        t = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device=dev)
        c = torch.tensor([float(nconf)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(c, op=dist.ReduceOp.SUM)
        return {"value": float(c) / (float(t) * 1e-3), "unit": "eloc_configs/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "steps": K, "ms_per_step": float(t) / K}


def ffma_peak_probe(dev):
    """FP32-FFMA throughput of this GPU, measured live (roofline denominator of the FFMA regime)."""
    import ctypes as C
    from rnnwavefunctions_b200 import _lib
    probe_out = torch.zeros(4, dtype=torch.float32, device=dev)
    best = 0.0
    blocks, iters = 148 * 4, 20000
    for _ in range(5):
        a, b = ev(), ev()
        a.record()
        _lib.lib.rnnwf_ffma_probe(C.c_void_p(probe_out.data_ptr()), blocks, iters, C.c_void_p(torch.cuda.current_stream().cuda_stream))
        b.record()
        torch.cuda.synchronize()
        best = max(best, blocks * 512 * iters * 16 * 2 / (a.elapsed_time(b) * 1e-3) / 1e12)
    return best


# the other configurations of BASELINE.json and the other precision regimes of the headline one, measured in the same run
# (N = 1 only; a few steps each) and reported under "workloads"
EXTRA_WORKLOADS = [("tfim1d_n20_gru50", "auto"), ("tfim1d_n20_gru50x3", "auto"), ("j1j2_n20_cgru50", "auto"), ("tfim2d_12x12_rnn50", "auto"),
                   ("heis10x10_sq_tgru100", "auto"), ("heis10x10_tri_tgru100", "tc3"), ("heis10x10_tri_tgru100", "fp32"),
                   ("rydberg_4x4_gru256", "auto"), ("rydberg_4x4_gru256", "tc3"), ("rydberg_8x8_gru256", "auto")]


def reference_arm(args, w, config, metric, unit, rank):
    """`--impl reference`: the reference's own implementation needs TensorFlow 1.13.1 (unavailable on this box), so the timed
    arm is the oracle port of its algorithm on all host cores, each step a bounded sample of the same workload, large enough
    that the chunked forwards carry full chunks of the reference's size (20 000 rows at C4)."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    # C4: 300 samples = ~40 000 configurations = two full 20 000-row chunks of the reference's chunk loop per step
    ns_ref = 300 if "heis" in args.workload else min(w["ns"], 200 if "rydberg" in args.workload else 500)
    for _ in range(max(0, min(args.warmup, 1))):
        oracle_step(w, 2)
    tot_c, tot_t = 0, 0.0
    for k in range(args.steps):
        c, dt, br = oracle_step(w, ns_ref, seed=k)
        tot_c += c
        tot_t += dt
    val = tot_c / tot_t
    line = {"impl": "reference", "metric": metric, "value": val, "unit": unit, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": val, "unit": unit, "cores": cores, "kind": "port",
                             "sample": f"{ns_ref} samples/step of the same workload ({tot_c // args.steps} configurations/step), oracle port "
                                       f"(oracle/: restated network forward + restated enumeration), torch float64, {cores} threads"},
            "e2e": {"value": val, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)

# ------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=list(WORKLOADS))
    ap.add_argument("--numsamples", type=int, default=None)
    ap.add_argument("--precision", default="auto", choices=["auto", "fp32", "tc", "tc3"],
                    help="sweep regime: fp32 = FFMA (1e-5 parity), tc = tcgen05 FP16-operand sweep (1e-3 bound), tc3 = tcgen05 "
                         "split-precision sweep (1e-5 parity); auto = the default regime of the cell (AUTO_PRECISION)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: numsamples per GPU fixed (default); strong: numsamples is the GLOBAL batch, sharded over the ranks")
    ap.add_argument("--no-prefix-reuse", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the other workloads / regimes reported under 'workloads' (N = 1)")
    args = ap.parse_args()
    w = dict(WORKLOADS[args.workload])
    if args.numsamples:
        w["ns"] = args.numsamples
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    metric, unit = "vmc_eloc_configs_per_s", "eloc_configs/s"
    ns_global = w["ns"] if args.scaling == "strong" else w["ns"] * world
    ns_rank = w["ns"] // world if args.scaling == "strong" else w["ns"]
    config = {"workload": f"{args.workload}: {w['desc']}", "numsamples_per_gpu": ns_rank, "numsamples_global": ns_global,
              "parallelism": f"dp{world} (samples sharded, weights replicated)",
              "l2": "per-step working set (saved hidden states + gates of the samples, ~2 GB) exceeds the 126 MB L2; no explicit flush"}

    if args.impl == "reference":
        reference_arm(args, w, config, metric, unit, rank)
        return

    import torch.distributed as dist
    if local_rank == 0:
        import __graft_entry__ as _g
        _g.build()                      # no-op when librnnwf.so is newer than its sources
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    B = Bench(args.workload, args.precision, ns=ns_rank, prefix_reuse=not args.no_prefix_reuse)
    precision = B.precision
    # rank 0 alone samples the clocks (of its own GPU); the sampler is set up (NVML initialised) before the warm-up steps
    clocks = ClockSampler(local_rank) if rank == 0 else None
    res = B.run(args.steps, args.warmup, barrier, clocks)
    clk = res["clocks"] or {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
    tmax = torch.tensor([res["elapsed_ms"]], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(res["nconf"]), float(ns_rank * args.steps)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    elapsed_ms = float(tmax)
    value = float(tot[0]) / (elapsed_ms * 1e-3)
    samples_per_s = float(tot[1]) / (elapsed_ms * 1e-3)

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    ffma_peak = ffma_peak_probe(dev)
    roofline = B.roofline(res, peaks, ffma_peak)
    if args.workload == DEFAULT_WORKLOAD and precision == "tc" and ns_rank == 10000:
        roofline["traffic"] = NCU_TRAFFIC["bytes"]
        roofline["traffic_source"] = NCU_TRAFFIC["source"]
    e2e = None if args.no_e2e else B.e2e(max(1, min(args.steps, 10)), barrier, world, dev)

    # strong scaling beside the weak number (N > 1, default mode): the same global batch as one GPU's, sharded over the ranks
    strong = None
    if world > 1 and args.scaling == "weak" and w["ns"] % world == 0:
        Bs = Bench(args.workload, precision, ns=w["ns"] // world, prefix_reuse=not args.no_prefix_reuse)
        rs = Bs.run(max(3, min(args.steps, 10)), 2, barrier)
        t2 = torch.tensor([rs["elapsed_ms"]], dtype=torch.float64, device=dev)
        c2 = torch.tensor([float(rs["nconf"])], dtype=torch.float64, device=dev)
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        dist.all_reduce(c2, op=dist.ReduceOp.SUM)
        strong = {"numsamples_global": w["ns"], "numsamples_per_gpu": w["ns"] // world, "ms_per_step": float(t2) / rs["steps"],
                  "value": float(c2) / (float(t2) * 1e-3), "unit": unit, "phase_ms_per_step": rs["phase_ms"],
                  "note": "same global batch as the N=1 run, sharded; efficiency = value / (N * value at N=1)"}
        del Bs

    if rank == 0:
        line = {"metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
                "dtype": DTYPE[precision], "precision_mode": precision, "data": "synthetic", "config": config,
                "samples_per_s": samples_per_s, "phase_ms_per_step": res["phase_ms"],
                "eloc_configs_per_step_per_gpu": res["nconf"] / args.steps, "gpu_launches": res["launches"],
                "clocks": clk, "roofline": roofline, "e2e": e2e, "engine": B.eng.describe(), "prefix_reuse": B.V.prefix_reuse}
        if strong is not None:
            line["strong_scaling"] = strong
        if world == 1 and not args.no_extra and args.workload == DEFAULT_WORKLOAD and args.numsamples is None:
            extra = {}
            del B
            torch.cuda.empty_cache()
            for name, prec in EXTRA_WORKLOADS:
                try:
                    Bx = Bench(name, prec)
                    small = "n20" in name                        # the sub-millisecond 1D configurations: more steps cost nothing
                    rx = Bx.run(30 if small else 3, 5 if small else 2, barrier)
                    rf = Bx.roofline(rx, peaks, ffma_peak)
                    ms = rx["elapsed_ms"] / rx["steps"]
                    extra[f"{name}:{Bx.precision}"] = {
                        "desc": Bx.w["desc"], "precision_mode": Bx.precision, "dtype": DTYPE[Bx.precision], "ms_per_step": ms,
                        "phase_ms_per_step": rx["phase_ms"], "eloc_configs_per_s": rx["nconf"] / (rx["elapsed_ms"] * 1e-3),
                        "samples_per_s": Bx.ns * rx["steps"] / (rx["elapsed_ms"] * 1e-3), "gpu_launches_per_step": rx["launches"] / rx["steps"],
                        "roofline": {k: rf[k] for k in ("bound", "kernel", "achieved", "peak", "unit", "frac")}, "steps": rx["steps"]}
                    del Bx
                    torch.cuda.empty_cache()
                except Exception as exc:       # one workload failing must not lose the headline line
                    extra[f"{name}:{prec}"] = {"error": f"{type(exc).__name__}: {exc}"}
            line["workloads"] = extra
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(w)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# dram__bytes_read.sum + dram__bytes_write.sum of one sweep_tc_kernel<MODE_CHILD> launch of the default workload, from the
# committed `ncu --set full` capture named here (a constant of that capture, not a measurement of this run)
NCU_TRAFFIC = {"bytes": 2.04e9, "source": "ncu --set full capture of commit 88ac7d1 (profiles/r2d_summary.md), one launch, 10^4 samples: dram__bytes_read 1.04 GB + dram__bytes_write 0.99 GB (17.1 GB before the ring slots were assigned by live range)"}


def executed_sites(w, samples, eng):
    """Sum over connected configurations of (sites - first changed generation step)."""
    e = w["eng"]
    nx, ny = e["nx"], e["ny"]
    T = nx * ny
    ham = w["ham"]
    ns = samples.shape[0]
    if ham in ("TFIM1D", "TFIM2D"):
        return ns * sum(T - t for t in range(T))
    if ham == "RYDBERG":
        return 3 * ns * sum(T - t for t in range(T))

    def step_of(ix, iy):
        return iy * nx + (ix if iy % 2 == 0 else nx - 1 - ix)

    total = 0
    if ham == "HEIS2D":
        J2 = w["hp"][0]
        s = samples.reshape(ns, nx, ny)
        for dx, dy, wgt in ((1, 0, 1.0), (0, 1, 1.0), (1, 1, J2)):
            if wgt == 0.0:
                continue
            for ix in range(nx - dx):
                for iy in range(ny - dy):
                    ts = min(step_of(ix, iy), step_of(ix + dx, iy + dy))
                    total += int((s[:, ix, iy] != s[:, ix + dx, iy + dy]).sum()) * (T - ts)
        return total
    if ham == "J1J2":
        N = nx
        hp = w["hp"]
        periodic = bool(hp[0])
        J1, J2 = hp[2:2 + N], hp[2 + N:2 + 2 * N]
        s = samples.reshape(ns, N)
        for d, J, lim in ((1, J1, N if periodic else N - 1), (2, J2, N if periodic else N - 2)):
            for i in range(lim):
                if J[i] != 0.0:
                    j = (i + d) % N
                    total += int((s[:, i] != s[:, j]).sum()) * (T - min(i, j))
        return total
    raise ValueError(ham)


if __name__ == "__main__":
    main()
