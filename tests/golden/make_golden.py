"""Generate the golden fixtures under ``tests/golden/`` by RUNNING THE UNMODIFIED REFERENCE.

Run in the build container only (``/root/reference`` does not exist on the GPU box):

    python tests/golden/make_golden.py

How the reference is run without TensorFlow / matplotlib (neither is installed here):

* ``Replicating_Portfolio.py`` is imported from ``/root/reference`` with stand-in modules for
  tensorflow, tqdm.notebook, matplotlib and seaborn.  Its functions first simulate every path
  array with NumPy/SciPy and only then build the Keras model; the stand-in for
  ``keras.initializers.RandomNormal`` (first Keras object constructed, line 149 / 350) raises,
  and the simulated arrays are read from the function's frame.  Nothing of the reference's
  simulation code is re-typed here.
* ``European Options.ipynb`` cells 3, 4, 6, 7, 8 are ``exec``-ed as stored in the notebook JSON,
  with the plotting calls landing on mocks.
* ``brownian_motion.py`` imports as is.

Outputs (all small): ``rp_paths.npz``, ``rp_sv_paths.npz``, ``eo_paths.npz``, ``brownian.npz``,
``sobol_points.npz``.
"""
from __future__ import annotations

import importlib.util
import json
import os
import sys
import types
import warnings
from math import ceil
from unittest import mock

import numpy as np

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


class _Captured(Exception):
    pass


def _install_stubs():
    def passthrough(f):
        return f

    tf = types.ModuleType("tensorflow")
    tf.function = passthrough
    keras = mock.MagicMock(name="keras")

    def _raise(*a, **k):
        raise _Captured()

    keras.initializers.RandomNormal = _raise
    tf.keras = keras
    tfk = types.ModuleType("tensorflow.keras")
    tfk.Model = mock.MagicMock()
    tfr = types.ModuleType("tensorflow.random")
    tfr.set_seed = lambda *_: None
    tf.random = tfr
    tqdm_nb = types.ModuleType("tqdm.notebook")
    tqdm_nb.tqdm = lambda x, *a, **k: x
    tqdm_nb.trange = lambda *a, **k: range(*a)
    mpl = types.ModuleType("matplotlib")
    plt = mock.MagicMock(name="pyplot")
    plt.subplots = lambda *a, **k: (mock.MagicMock(), (mock.MagicMock(), mock.MagicMock()))
    mpl.pyplot = plt
    mpl.ticker = mock.MagicMock()
    sys.modules.update({
        "tensorflow": tf, "tensorflow.keras": tfk, "tensorflow.random": tfr,
        "tqdm.notebook": tqdm_nb, "matplotlib": mpl, "matplotlib.pyplot": plt,
        "matplotlib.ticker": mpl.ticker, "seaborn": mock.MagicMock(),
    })
    return plt


def _load_reference_module(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _run_and_capture(fn, params):
    try:
        fn(params)
    except _Captured:
        tb = sys.exc_info()[2]
        while tb.tb_next is not None:
            tb = tb.tb_next
        frame = tb.tb_frame
        # innermost frame is the stub; walk up to the reference function
        while frame is not None and frame.f_code.co_name != fn.__name__:
            frame = frame.f_back
        return dict(frame.f_locals)
    raise RuntimeError("reference ran past the Keras construction?")


def golden_rp(RP):
    mu, sigma_0 = 0.09464, 0.15965          # Multi Time Step.ipynb cell 9 / 28
    params = {'Y': 1, 'K': 1, 'T': 10, 'mu': mu, 'r': 0.03, 'sigma': sigma_0, 'rebalancing': 1 / 4,
              'N': 10_000, 'P': 100, 'x': 55.0, 'l0': 0.01, 'c': 0.075, 'ita': 0.000597,
              'dt': 1 / 100, 'n_paths': 7}
    loc = _run_and_capture(RP.Replicating_Portfolio, params)
    Yfull = loc['Y_paths'].base if loc['Y_paths'].base is not None else loc['Y_paths']
    np.savez_compressed(
        os.path.join(OUT, "rp_paths.npz"),
        params=json.dumps(params), Y_reduced=loc['Y_paths'], Y_full_cols=Yfull[:, ::125],
        Y_full_mean_T=Yfull[:, -1].mean(), B_reduced=np.array(loc['B'][0]),
        L_cols=loc['L_paths'][:, ::125], N_T=loc['N_paths'][:, -1],
        Payoff_Y=loc['Payoff_Y'], out_of_money_P=loc['out_of_money_P'],
        reduction=loc['reduction'], n_time_steps=loc['n_time_steps'], dt=loc['dt'])
    # the notebook-size run: only summary statistics are kept
    params_big = dict(params, n_paths=ceil(np.log2(3_000)))
    loc = _run_and_capture(RP.Replicating_Portfolio, params_big)
    return {"rp_big_mean_Y_T": float(loc['Y_paths'][:, -1].mean()),
            "rp_big_out_of_money_P": float(loc['out_of_money_P']),
            "rp_big_mean_payoff": float(loc['Payoff_Y'].mean())}


def golden_rp_sv(RP):
    # Multi Time Step.ipynb cell 32 — the dict literal has two 'c' keys; the last one wins.
    params = {'Y': 1, 'K': 1, 'T': 10, 'mu': 0.09620, 'r': 0.03, 's0': 0.16679, 'rebalancing': 1 / 4,
              'a': 0.0033281299103885727, 'b': 0.1562947229160206, 'c': 0.01583113545983743,
              'N': 10_000, 'P': 100, 'x': 55.0, 'l0': 0.01, 'c': 0.075, 'ita': 0.000597,  # noqa: F601
              'dt': 1 / 100, 'n_paths': 7}
    out = {}
    for tag, c in (("c075", 0.075), ("c0158", 0.01583113545983743)):
        p = dict(params, c=c)
        loc = _run_and_capture(RP.Replicating_Portfolio_SV, p)
        Yfull = loc['Y_paths'].base if loc['Y_paths'].base is not None else loc['Y_paths']
        out[tag] = dict(Y_reduced=loc['Y_paths'], Y_full_cols=Yfull[:, ::125], vt_cols=loc['vt'][:, ::125],
                        Payoff_Y=loc['Payoff_Y'], out_of_money_P=loc['out_of_money_P'])
    np.savez_compressed(
        os.path.join(OUT, "rp_sv_paths.npz"), params=json.dumps(params),
        **{f"{tag}_{k}": v for tag, d in out.items() for k, v in d.items()})


def golden_eo(plt):
    import pandas as pd
    nb = json.load(open(os.path.join(REF, "European Options.ipynb")))
    cells = ["".join(c["source"]) for c in nb["cells"]]
    summary = {}
    with mock.patch.object(pd.DataFrame, "plot", mock.MagicMock()), \
            mock.patch.object(pd.Series, "hist", mock.MagicMock()):
        for tag, n_paths_override in (("nb", None), ("small", 100)):
            g = {"__name__": "eo"}
            exec("import numpy as np, pandas as pd\nfrom math import log, sqrt, exp, erf, ceil, floor\n"
                 "import scipy.stats as stats\nfrom scipy.stats import qmc\nfrom time import perf_counter\n", g)
            g["plt"] = plt
            exec(cells[3], g)
            if n_paths_override:
                g["N_paths"] = n_paths_override
            exec(cells[4], g)
            exec(cells[6], g)
            S_fine_T_mean = float(np.nanmean(g["S_paths"][:, -1]))
            S_fine = g["S_paths"]
            exec(cells[7], g)
            exec(cells[8], g)
            if tag == "nb":
                summary.update(eo_mean_S_T_fine=S_fine_T_mean,
                               eo_mean_S_T_weekly=float(g["S_paths"][:, -1].mean()),
                               eo_payoff_mean=float(g["payoff"].mean()),
                               eo_payoff_std=float(g["payoff"].std(ddof=1)),
                               eo_payoff_max=float(g["payoff"].max()),
                               eo_prob_otm=float(g["prob_out_of_money"]))
            else:
                np.savez_compressed(
                    os.path.join(OUT, "eo_paths.npz"),
                    S0=g["S0"], K=g["K"], r=g["r"], sigma=g["sigma"], T=g["T"], dt_fine=1 / 365,
                    n_paths_log2=g["n_paths"], S_fine_cols=S_fine[:, ::73], S_weekly=g["S_paths"],
                    B_weekly=np.array(g["B"][0]), payoff=g["payoff"], prob_otm=g["prob_out_of_money"],
                    reduction=g["reduction"], n_time_steps=g["n_time_steps"])
    return summary


def golden_brownian():
    bm = _load_reference_module("brownian_motion")
    np.savez_compressed(os.path.join(OUT, "brownian.npz"),
                        dW_5_1=bm.get_dW(5, 1), W_5_1=bm.get_W(5, 1),
                        dW_64_1234=bm.get_dW(64, 1234), W_64_1234=bm.get_W(64, 1234))


def golden_sobol():
    """Integer points straight from SciPy for the five BASELINE shapes (windows only)."""
    from scipy.stats import qmc
    out = {}
    for d, seed, m in ((2, 1235, 16), (53, 1235, 20), (253, 1235, 22), (253, 1234, 22), (253, 1235, 26)):
        eng = qmc.Sobol(d, scramble=True, seed=seed)
        out[f"sv_{d}_{seed}"] = eng._sv.copy()
        out[f"shift_{d}_{seed}"] = eng._shift.copy()
        for start in (0, 2 ** m - 64, (2 ** m) // 3):
            e = qmc.Sobol(d, scramble=True, seed=seed)
            if start:
                e.fast_forward(start)
            cols = sorted({0, 1, d // 2, d - 1})
            out[f"pts_{d}_{seed}_{start}"] = (e.random(64) * 2.0 ** 30).astype(np.uint32)[:, cols]
            out[f"cols_{d}_{seed}_{start}"] = np.array(cols)
    np.savez_compressed(os.path.join(OUT, "sobol_points.npz"), **out)


def notebook_fit_log():
    """What the reference's own Keras run of ``European Options.ipynb`` left in the saved notebook: the training log
    of cell 13 (52 ``model1.fit`` calls on 4096 paths, batch 512) and the prints of cells 18 and 20.  TensorFlow is
    not installed, so these stored outputs are the only evidence of the reference's fit; the fit tests hold the
    oracle and the CUDA path against them statistically (``tests/test_oracle_fit.py``, ``tests/test_gpu_fit.py``)."""
    import re
    with open(os.path.join(REF, "European Options.ipynb")) as f:
        cells = json.load(f)["cells"]

    def text(i):
        return "".join("".join(o.get("text", [])) for o in cells[i].get("outputs", []) if "text" in o)

    log = text(13)
    ep = re.findall(r"Epoch (\d+)/(\d+)", log)
    starts = [i for i, (a, _) in enumerate(ep) if a == "1"]
    runs = [int(ep[(starts[k + 1] - 1) if k + 1 < len(starts) else -1][0]) for k in range(len(starts))]
    # the last progress line of every epoch carries that epoch's mean loss
    per_epoch = re.findall(r"8/8 \[=+\] - [^\n]*? loss: ([0-9.e+-]+) - mae: ([0-9.e+-]+)", log)
    var = [[float(a), float(b)] for a, b in re.findall(r"VaR: ([0-9.]+) \(98%\),\s+([0-9.]+) \(99%\)", log)]
    out18, out20 = text(18), text(20)
    # every fit is followed by one model.evaluate line in the same format: epochs + 1 entries per fit
    idx, fit_last_loss, eval_loss = 0, [], []
    for n in runs:
        fit_last_loss.append(float(per_epoch[idx + n - 1][0]))
        eval_loss.append(float(per_epoch[idx + n][0]))
        idx += n + 1
    assert idx == len(per_epoch), (idx, len(per_epoch))
    return {
        "paths": 4096, "columns": 53, "params": {"S0": 100, "K": 100, "r": 0.08, "sigma": 0.15, "T": 1.0, "dt": "1/365",
                                                "rebalancing": "1/52"},
        "epochs_per_fit": runs, "fit_last_epoch_loss": fit_last_loss, "evaluate_loss": eval_loss, "VaR_98_99": var,
        "phi0": float(re.search(r"Phi t=0 : ([0-9.]+)", out18).group(1)),
        "psi0": float(re.search(r"Psi t=0 : ([0-9.]+)", out18).group(1)),
        "V0_RP": float(re.search(r"Value at t_0 \(Replicating-P\)\s+= ([0-9.]+)", out20).group(1)),
        # cell 20 labels this "Discounted Eq[V(T)]" but computes values[:, 0].mean() * exp(-r T): V0_RP discounted once more
        "printed_discounted_V0": float(re.search(r"Discounted Eq\[V\(T\)\]\s+= ([0-9.]+)", out20).group(1)),
    }


def notebook_rp_log():
    """Saved output of ``RP.Replicating_Portfolio(parameters)`` and ``RP.Replicating_Portfolio_SV(sv_parameters)`` in
    ``Multi Time Step.ipynb`` (cells 28 and 32): the ``>> Y_(t) = ..., N_(t) = ...`` line of every rebalancing step
    (means of the simulated fund and survivor columns: path side, reproducible to the printed digits), the
    ``model.evaluate`` loss and the VaR line of every step, and the returned hedge weights (fit side: statistical)."""
    import re
    with open(os.path.join(REF, "Multi Time Step.ipynb")) as f:
        cells = json.load(f)["cells"]
    out = {}
    for tag, i in (("rp", 28), ("rp_sv", 32)):
        txt = "".join("".join(o.get("text", [])) for o in cells[i].get("outputs", []) if "text" in o)
        out[tag] = {
            "step_lines": re.findall(r">> Y_\([0-9.]+\) = [0-9.]+, N_\([0-9.]+\) = [0-9.]+", txt),
            "evaluate_loss": [float(x) for x in re.findall(r"8/8 \[=+\] - [^\n]*? loss: ([0-9.e+-]+) - mae", txt)],
            "VaR_98_99": [[float(a), float(b)] for a, b in re.findall(r"VaR: (-?[0-9.]+) \(98%\),\s+(-?[0-9.]+) \(99%\)", txt)],
            "phi0": int(re.search(r"Phi t=0 : ([0-9,]+)", txt).group(1).replace(",", "")),
            "psi0": int(re.search(r"Psi t=0 : ([0-9,]+)", txt).group(1).replace(",", "")),
        }
    # cell 29: the same call for five volatilities; cell 30 shows the returned weights as a table
    txt = "".join("".join(o.get("text", [])) for o in cells[29].get("outputs", []) if "text" in o)
    parts = re.split(r">>(0\.\d+)-+", txt)
    table = "".join(cells[30]["outputs"][0]["data"]["text/plain"])
    rows = {float(a): (float(b), float(c)) for a, b, c in re.findall(r"^(0\.\d+)\s+([0-9.]+)\s+([0-9.]+)\s", table, re.M)}
    out["sigma_sweep"] = {
        sig: {"step_lines": re.findall(r">> Y_\([0-9.]+\) = [0-9.]+, N_\([0-9.]+\) = [0-9.]+", body),
              "phi0": rows[float(sig)][0], "psi0": rows[float(sig)][1]}
        for sig, body in zip(parts[1::2], parts[2::2])}
    return out


def main():
    warnings.filterwarnings("ignore")
    plt = _install_stubs()
    RP = _load_reference_module("Replicating_Portfolio")
    summary = {}
    summary.update(golden_rp(RP))
    golden_rp_sv(RP)
    summary.update(golden_eo(plt))
    golden_brownian()
    golden_sobol()
    # numbers printed in the reference's saved notebook outputs
    summary["notebook_printed"] = {
        "STS_mean_Y_T": 2.2201138207241464, "EO_mean_S_T_fine": 108.327487,
        "EO_mean_S_T_weekly": 108.30487063199344, "EO_payoff_mean": 11.258669,
        "EO_payoff_std": 12.972705, "EO_payoff_max": 100.871992, "EO_prob_otm": 0.3293,
        "MTS_mean_Y_T": 2.227189, "brownian_get_W_5_1": [0, 1.62434536, 1.01258895, 0.4844172, -0.58855142],
    }
    summary["notebook_fit_log"] = notebook_fit_log()
    summary["notebook_rp_log"] = notebook_rp_log()
    with open(os.path.join(OUT, "summary.json"), "w") as f:
        json.dump(summary, f, indent=1, sort_keys=True)
    print(json.dumps(summary, indent=1, sort_keys=True))


if __name__ == "__main__":
    main()
