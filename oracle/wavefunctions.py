"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py) -- float64 restatement of the reference's
wavefunction classes.  Everything is plain ``torch`` float64 on the CPU so that autograd of the
reference's surrogate cost gives the oracle gradient the same way TF autodiff does upstream
(``optimizer.compute_gradients(cost)``, 1DTFIM/TrainingRNN_1DTFIM.py:161).

Each function takes a dict ``P`` of named parameter tensors in the *TensorFlow layouts*:

  1D GRU (tf.contrib.cudnn_rnn.CudnnCompatibleGRUCell, TF 1.13.1, not vendored)
      gru{l}.Wg  [d_in+nh, 2nh]  rows [x;h], cols [r|u]      gru{l}.bg  [2nh]
      gru{l}.Wci [d_in, nh]   gru{l}.Wch [nh, nh]   gru{l}.bci [nh]   gru{l}.bch [nh]
  2D vanilla (2DTFIM_2DRNN/MDRNNcell.py:21-35)  Wh [nh,nh] Uh [2,nh] Wv [nh,nh] Uv [2,nh] b [nh]
  tensorized (2DHeisenberg/MDTensorized{,GRU}cell.py:20-33) W,Wg [nh,2nh,4]  b,bg [nh]  Wmerge [2nh,nh]
  Rydberg GRU (RydbergKagome_2DGRU/MDGRUcell.py:19-32) W,Wg [4(nh+6),nh] b,bg [nh] Wmerge [4nh,nh]
  heads  dense.W [nh,D] dense.b [D]  (+ dense_phase.W/.b for the complex models)

Sampling is ancestral with *explicit uniforms* ``u[b, t]`` (t = generation-order step) and the
inverse-CDF rule of TF's CPU multinomial kernel: idx = #{j : cdf_j <= u * total}.
"""
import math

import torch

F64 = torch.float64


# --------------------------------------------------------------------------------------------
# helpers
# --------------------------------------------------------------------------------------------
def heavyside(x):
    """Step with H(0)=1 -- J1J2/ComplexRNNwavefunction.py:11-13, 2DHeisenberg/MDComplexRNN.py:19-21."""
    return (x >= 0).to(F64)


def inverse_cdf_draw(p, u):
    """tf.multinomial / tf.random.categorical CPU kernel semantics.

    p [B,D] non-negative (masked classes are exactly 0 and are skipped like log(0)=-inf entries),
    u [B] in [0,1).  Returns (idx [B] int64, margin [B]) where margin is the distance of the draw
    to the nearest CDF threshold relative to the total (used by tests for the 1e-6 rule).
    """
    cdf = torch.cumsum(p, dim=1)
    total = cdf[:, -1:]
    target = u.reshape(-1, 1).to(F64) * total
    idx = (cdf <= target).sum(dim=1)
    idx = torch.clamp(idx, max=p.shape[1] - 1)
    margin = ((cdf[:, :-1] - target).abs() / total).min(dim=1).values
    return idx, margin


def one_hot(idx, depth):
    return torch.nn.functional.one_hot(idx.long(), depth).to(F64)


def u1_mask(p, n_up, n_generated, N, magnetization=0):
    """U(1) projection used by both complex models.

    2DHeisenberg/MDComplexRNN.py:69-76 (L1 renormalisation of probabilities, every site) and
    J1J2/ComplexRNNwavefunction.py:85-93,147-155 (L2 renormalisation of sqrt-probabilities for
    n >= N/2; identical to the L1 rule on probabilities, and a no-op for n < N/2)."""
    n_dn = n_generated - n_up
    act_up = heavyside(((N + magnetization) // 2 - 1) - n_up)
    act_dn = heavyside(((N - magnetization) // 2 - 1) - n_dn)
    p = p * torch.stack([act_dn, act_up], dim=1)
    return p / p.sum(dim=1, keepdim=True)


def zigzag_order(Nx, Ny):
    """Generation order of the 2D models: outer ny, inner nx left->right on even rows and
    right->left on odd rows (2DTFIM_2DRNN/RNNwavefunction.py:90-113)."""
    order = []
    for ny in range(Ny):
        xs = range(Nx) if ny % 2 == 0 else range(Nx - 1, -1, -1)
        for nx in xs:
            order.append((nx, ny))
    return order


# --------------------------------------------------------------------------------------------
# 1D GRU stack  (1DTFIM/RNNwavefunction.py, J1J2/ComplexRNNwavefunction.py)
# --------------------------------------------------------------------------------------------
def gru_cell(P, l, x, h):
    """CudnnCompatibleGRUCell.call (TF 1.13.1): reset gate applied after the hidden matmul."""
    nh = h.shape[1]
    gates = torch.sigmoid(torch.cat([x, h], dim=1) @ P[f"gru{l}.Wg"] + P[f"gru{l}.bg"])
    r, u = gates[:, :nh], gates[:, nh:]
    c = torch.tanh(x @ P[f"gru{l}.Wci"] + P[f"gru{l}.bci"] + r * (h @ P[f"gru{l}.Wch"] + P[f"gru{l}.bch"]))
    return (1.0 - u) * c + u * h


def rnn1d(P, N, units, samples=None, uniforms=None, numsamples=None, complex_=False, u1=False):
    """1D pRNN / cRNN sweep.

    pRNN:  1DTFIM/RNNwavefunction.py:35-74 (sample), :76-118 (log_probability).
    cRNN:  J1J2/ComplexRNNwavefunction.py:45-103 (sample), :105-169 (log_amplitude).
    Returns dict(samples [B,N] int64, logp [B], phase [B], margin [B,N]).
    For the cRNN  log psi = 0.5*logp + i*phase.
    """
    sampling = samples is None
    B = numsamples if sampling else samples.shape[0]
    L = len(units)
    h = [torch.zeros(B, units[l], dtype=F64) for l in range(L)]
    x = torch.zeros(B, 2, dtype=F64)  # first input is the zero vector, not a one-hot (:52-55)
    logp = torch.zeros(B, dtype=F64)
    phase = torch.zeros(B, dtype=F64)
    n_up = torch.zeros(B, dtype=F64)
    out, margins = [], []
    for n in range(N):
        inp = x
        for l in range(L):
            h[l] = gru_cell(P, l, inp, h[l])
            inp = h[l]
        p = torch.softmax(inp @ P["dense.W"] + P["dense.b"], dim=1)
        if u1:
            p = u1_mask(p, n_up, float(n), N, 0)
        if sampling:
            s, m = inverse_cdf_draw(p.detach(), uniforms[:, n])
            margins.append(m)
        else:
            s = samples[:, n].long()
        out.append(s)
        logp = logp + torch.log(p.gather(1, s.reshape(-1, 1)).squeeze(1))
        if complex_:
            ph = math.pi * torch.nn.functional.softsign(inp @ P["dense_phase.W"] + P["dense_phase.b"])
            phase = phase + ph.gather(1, s.reshape(-1, 1)).squeeze(1)
        n_up = n_up + s.to(F64)
        x = one_hot(s, 2)
    res = dict(samples=torch.stack(out, dim=1), logp=logp, phase=phase)
    if sampling:
        res["margin"] = torch.stack(margins, dim=1)
    return res


def rnn1d_paritysym_logp(P, N, units, samples):
    """1DTFIM/RNNwavefunction_paritysym.py:122-145: log(0.5*(P(s) + P(reversed s)))."""
    lp1 = rnn1d(P, N, units, samples=samples)["logp"]
    lp2 = rnn1d(P, N, units, samples=torch.flip(samples, dims=[1]))["logp"]
    return torch.log(0.5 * (torch.exp(lp1) + torch.exp(lp2)))


# --------------------------------------------------------------------------------------------
# 2D cells
# --------------------------------------------------------------------------------------------
def vanilla2d_cell(P, xs, hs):
    """2DTFIM_2DRNN/MDRNNcell.py:51-66."""
    pre = xs[0] @ P["Uh"] + hs[0] @ P["Wh"] + xs[1] @ P["Uv"] + hs[1] @ P["Wv"] + P["b"]
    return torch.nn.functional.elu(pre)


def tensorized_cell(P, xs, hs, gru):
    """2DHeisenberg/MDTensorizedcell.py:40-49, MDTensorizedGRUcell.py:49-61.
    tensordot(T, W, axes=[[1,2],[1,2]]) with T = hc (x) xc  ==  einsum('bik,jik->bj')."""
    hc = torch.cat(hs, dim=1)
    xc = torch.cat(xs, dim=1)
    T = torch.einsum("bi,bk->bik", hc, xc)
    s = torch.einsum("bik,jik->bj", T, P["W"])
    if not gru:
        return torch.tanh(s + P["b"])
    g = torch.einsum("bik,jik->bj", T, P["Wg"])
    u = torch.sigmoid(g + P["bg"])
    ht = torch.tanh(s + P["b"])
    return u * ht + (1.0 - u) * (hc @ P["Wmerge"])


def rydberg_cell(P, xs, hs):
    """RydbergKagome_2DGRU/MDGRUcell.py:48-59."""
    z = torch.cat(list(xs) + list(hs), dim=1)
    ht = torch.tanh(z @ P["W"] + P["b"])
    u = torch.sigmoid(z @ P["Wg"] + P["bg"])
    return u * ht + (1.0 - u) * (torch.cat(list(hs), dim=1) @ P["Wmerge"])


def rnn2d(P, Nx, Ny, nh, cell, samples=None, uniforms=None, numsamples=None,
          complex_=False, mag_fixed=False, magnetization=0):
    """2D open-boundary sweep with two neighbours.

    cell in {"vanilla", "tens", "tensgru"}.
    2DTFIM_2DRNN/RNNwavefunction.py:35-118 / :120-200 ;  2DHeisenberg/MDComplexRNN.py:80-173 / :177-286.
    Configurations are indexed [b, nx, ny].  Returns dict(samples, logp, phase[, margin]).
    """
    sampling = samples is None
    B = numsamples if sampling else samples.shape[0]
    N = Nx * Ny
    zero_h = torch.zeros(B, nh, dtype=F64)
    zero_x = torch.zeros(B, 2, dtype=F64)
    H, X = {}, {}
    logp = torch.zeros(B, dtype=F64)
    phase = torch.zeros(B, dtype=F64)
    n_up = torch.zeros(B, dtype=F64)
    out = torch.zeros(B, Nx, Ny, dtype=torch.int64)
    margins = []
    for t, (nx, ny) in enumerate(zigzag_order(Nx, Ny)):
        px = nx - 1 if ny % 2 == 0 else nx + 1  # in-row predecessor
        hs = (H.get((px, ny), zero_h), H.get((nx, ny - 1), zero_h))
        xs = (X.get((px, ny), zero_x), X.get((nx, ny - 1), zero_x))
        if cell == "vanilla":
            h = vanilla2d_cell(P, xs, hs)
        else:
            h = tensorized_cell(P, xs, hs, gru=(cell == "tensgru"))
        H[(nx, ny)] = h
        p = torch.softmax(h @ P["dense.W"] + P["dense.b"], dim=1)
        if mag_fixed:
            p = u1_mask(p, n_up, float(t), N, magnetization)
        if sampling:
            s, m = inverse_cdf_draw(p.detach(), uniforms[:, t])
            margins.append(m)
        else:
            s = samples[:, nx, ny].long()
        out[:, nx, ny] = s
        logp = logp + torch.log(p.gather(1, s.reshape(-1, 1)).squeeze(1))
        if complex_:
            ph = math.pi * torch.nn.functional.softsign(h @ P["dense_phase.W"] + P["dense_phase.b"])
            phase = phase + ph.gather(1, s.reshape(-1, 1)).squeeze(1)
        n_up = n_up + s.to(F64)
        X[(nx, ny)] = one_hot(s, 2)
    res = dict(samples=out, logp=logp, phase=phase)
    if sampling:
        res["margin"] = torch.stack(margins, dim=1)
    return res


def rydberg2d(P, Lx, Ly, nh, samples=None, uniforms=None, numsamples=None):
    """Periodic 2D GRU over Kagome unit cells, 8-way softmax per cell.

    RydbergKagome_2DGRU/MDRNNWavefunction.py:73-153 (sample), :157-238 (log_probability).
    samples [B, Lx, Ly, 3] bits; class c = b0 + 2 b1 + 4 b2 (:12-34).
    Neighbour order fed to the cell (:117-119,136-138): in-row predecessor, (nx,ny-1), in-row successor
    mod Lx, (nx,(ny+1) mod Ly); a neighbour that has not been generated yet contributes zeros
    (intended lattice semantics; SURVEY App. C-7 on the dict-key collisions of the reference).
    """
    sampling = samples is None
    B = numsamples if sampling else samples.shape[0]
    zero_h = torch.zeros(B, nh, dtype=F64)
    zero_x = torch.zeros(B, 6, dtype=F64)
    H, X = {}, {}
    logp = torch.zeros(B, dtype=F64)
    out = torch.zeros(B, Lx, Ly, 3, dtype=torch.int64)
    margins = []
    for t, (nx, ny) in enumerate(zigzag_order(Lx, Ly)):
        if ny % 2 == 0:
            keys = [(nx - 1, ny), (nx, ny - 1), ((nx + 1) % Lx, ny), (nx, (ny + 1) % Ly)]
        else:
            keys = [(nx + 1, ny), (nx, ny - 1), ((nx - 1) % Lx, ny), (nx, (ny + 1) % Ly)]
        hs = [H.get(k, zero_h) for k in keys]
        xs = [X.get(k, zero_x) for k in keys]
        h = rydberg_cell(P, xs, hs)
        p = torch.softmax(h @ P["dense.W"] + P["dense.b"], dim=1)
        if sampling:
            c, m = inverse_cdf_draw(p.detach(), uniforms[:, t])
            margins.append(m)
            bits = torch.stack([c % 2, (c // 2) % 2, (c // 4) % 2], dim=1)
        else:
            bits = samples[:, nx, ny, :].long()
            c = bits[:, 0] + 2 * bits[:, 1] + 4 * bits[:, 2]
        out[:, nx, ny, :] = bits
        logp = logp + torch.log(p.gather(1, c.reshape(-1, 1)).squeeze(1))
        H[(nx, ny)] = h
        X[(nx, ny)] = one_hot(bits, 2).reshape(B, 6)
    res = dict(samples=out, logp=logp, phase=torch.zeros(B, dtype=F64))
    if sampling:
        res["margin"] = torch.stack(margins, dim=1)
    return res


# --------------------------------------------------------------------------------------------
# symmetry-projected amplitudes (P2 rows W15 / W19 of SURVEY 8a)
# --------------------------------------------------------------------------------------------
def regularized_identity(x, eps=1e-4):
    """2DHeisenberg/MDComplexRNN.py:23-25 (sign with sign(0)=+1, stop-gradient on the sign)."""
    sign = torch.where(x >= 0, torch.ones_like(x), -torch.ones_like(x)).detach()
    return sign * torch.sqrt(x ** 2 + eps ** 2)


def log_amplitudes_from_symmetry_group(logp_g, phase_g, characters):
    """2DHeisenberg/MDComplexRNN.py:290-316. logp_g, phase_g: [G,B]; characters: list of +-1.
    Returns complex log-amplitude [B].  As upstream (:313), the phase sum runs over range(len(characters)) -- normally G;
    see `heis_symmetry_copies(legacy_list_bug=True)` for the one case where the reference passes fewer."""
    G = logp_g.shape[0]
    Gc = len(characters)
    ch = torch.as_tensor(characters, dtype=F64).reshape(Gc, 1)
    logp = torch.log(torch.exp(logp_g).sum(0) / G)
    re = (torch.cos(phase_g[:Gc]) / ch).sum(0)
    im = (torch.sin(phase_g[:Gc]) / ch).sum(0)
    ph = torch.atan2(im, regularized_identity(re))
    return torch.complex(0.5 * logp, ph)


def rotate60(samples):
    """RydbergKagome_2DGRU/MDRNNWavefunction.py:281-291."""
    B, Lx, Ly, _ = samples.shape
    out = torch.zeros_like(samples)
    for i in range(Lx):
        for j in range(Ly):
            out[:, i, j, 0] = samples[:, (i + j + 1) % Lx, (Ly - i) % Ly, 1]
            out[:, i, j, 1] = samples[:, (i + j) % Lx, (Ly - i) % Ly, 2]
            out[:, i, j, 2] = samples[:, (i + j + 1) % Lx, (Ly - (i + 1)) % Ly, 0]
    return out


# group elements in the order the reference appends them to list_samples, and the character tables
# (2DHeisenberg/MDComplexRNN.py: rotsym :320-333, rotsym_spinparity :335-356, rotreducedsym(_spinparity) :358-389,
#  c2vsym :391-408, c2v_spinparity :410-435, c2sym(_spinparity) :438-469, c2dsym(_spinparity) :471-518,
#  c4vsym(_spinparity) :520-577).  tf.image.rot90(k) on [b, Nx, Ny, 1] == np.rot90(k, axes=(1, 2)).
def _rot(k):
    return lambda s: torch.rot90(s, k, dims=(1, 2))


HEIS_GROUP_OPS = {
    "id": lambda s: s, "r1": _rot(-1), "r2": _rot(-2), "r3": _rot(-3),
    "fx": lambda s: torch.flip(s, dims=[1]), "fy": lambda s: torch.flip(s, dims=[2]),
    "t": lambda s: s.transpose(1, 2), "tr2": lambda s: torch.rot90(s, -2, dims=(1, 2)).transpose(1, 2),
}
_C2V_TABLE = {"A1": [1, 1, 1, 1], "A2": [1, 1, -1, -1], "B1": [1, -1, 1, -1], "B2": [1, -1, -1, 1]}
HEIS_GROUPS = {
    "rotsym": (["id", "r1", "r2", "r3"], {"A": [1, 1, 1, 1], "B": [1, -1, 1, -1]}),
    "rotreducedsym": (["id", "r1"], {"A": [1, 1], "B": [1, -1]}),
    "c2sym": (["id", "r2"], {"A": [1, 1], "B": [1, -1]}),
    "c2vsym": (["id", "r2", "fx", "fy"], _C2V_TABLE),
    "c2dsym": (["id", "r2", "t", "tr2"], _C2V_TABLE),
    "c4vsym": (["id", "r1", "r2", "r3", "fx", "fy", "t", "tr2"],
               {"A1": [1] * 8, "A2": [1, 1, 1, 1, -1, -1, -1, -1], "B1": [1, -1, 1, -1, 1, 1, -1, -1],
                "B2": [1, -1, 1, -1, -1, -1, 1, 1]}),
}


def heis_symmetry_copies(samples, group, group_character, spinparity_value=None, legacy_list_bug=False):
    """-> (list of transformed configurations [B,Nx,Ny], list of character signs).  With a spin parity the spin-flipped
    copies 1 - s follow in the same order with signs spinparity_value * chi (e.g. :343-346, :354).

    Reference defect reproduced only on request (`legacy_list_bug=True`, used to check the fixtures): every spin-parity
    variant except c4vsym keeps chi as a Python *list*, so `spinparity_value*group_character_signs` (:354, :373, :433,
    :453, :516) is list repetition -- for spinparity_value = -1 it is the empty list, and the phase sum of
    log_amplitudes_fromsymmetrygroup then runs over the un-flipped copies only (the amplitude still averages all 2G).
    c4vsym_spinparity uses np.array (:566-575) and multiplies as intended.  The product implements the intended signs."""
    ops, table = HEIS_GROUPS[group]
    lst = [HEIS_GROUP_OPS[o](samples) for o in ops]
    signs = list(table[group_character])
    if spinparity_value is not None:
        lst += [HEIS_GROUP_OPS[o](1 - samples) for o in ops]
        if legacy_list_bug and group != "c4vsym":
            signs = signs + int(spinparity_value) * signs          # list repetition, as upstream
        else:
            signs = signs + [spinparity_value * c for c in signs]
    return lst, signs


def rydberg_symmetry_copies(samples, symmetry):
    """RydbergKagome_2DGRU/MDRNNWavefunction.py:246-263: the six 60-degree rotations, or all Lx*Ly translations
    (tf.roll by (rx, ry) along the cell axes, rx outer)."""
    if symmetry == "c6sym":
        lst = [samples]
        for _ in range(5):
            lst.append(rotate60(lst[-1]))
        return lst
    if symmetry == "translationsym":
        _, Lx, Ly, _ = samples.shape
        return [torch.roll(samples, shifts=(rx, ry), dims=(1, 2)) for rx in range(Lx) for ry in range(Ly)]
    raise ValueError(symmetry)


def log_probability_symmetry(logp_g):
    """RydbergKagome_2DGRU/MDRNNWavefunction.py:275-279: log(mean_g P_g); logp_g [G,B]."""
    return torch.log(torch.exp(logp_g).sum(0) / logp_g.shape[0])
