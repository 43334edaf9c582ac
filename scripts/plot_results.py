#!/usr/bin/env python
"""Result and latent plots (the reference's ``src/Visualization/*.ipynb`` as a script).

  python scripts/plot_results.py results  <output_root> [--out auc.pdf]
  python scripts/plot_results.py latents  <latent_*.pkl> [--round -2] [--out latent.pdf]

``results`` reads every ``*_results.json`` JSON-lines file under ``Checkpoint/Results`` and draws mean
client AUC per round for each (model_type, update_type).  ``latents`` reads a latent dump (format:
``dict[round][device] -> (latent [N,7], label [N])``, the contract of
``src/Visualization/latent_visualization.ipynb:40-59``), takes ``rounds[--round]`` like the notebook
and draws a 2-D t-SNE (or PCA when scikit-learn's TSNE is unavailable) per device.
matplotlib is optional: without it the script prints the tables instead.
"""
import argparse
import glob
import json
import os
import pickle
import sys

import numpy as np


def load_results(root):
    out = {}
    for path in glob.glob(os.path.join(root, "Checkpoint/Results/**/*_results.json"), recursive=True):
        rows = [json.loads(l) for l in open(path) if l.strip()]
        if rows and "model_type" in rows[0]:      # skips verification_results.json
            key = (rows[0]["model_type"], rows[0]["update_type"], os.path.relpath(path, root))
            out[key] = [(r["round"], float(np.mean(r["client_metrics"])), r["global_loss"]) for r in rows]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["results", "latents"])
    ap.add_argument("path")
    ap.add_argument("--round", type=int, default=-2)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except Exception:
        plt = None
    if a.what == "results":
        res = load_results(a.path)
        for (mt, ut, p), rows in sorted(res.items()):
            print(f"{mt:12s} {ut:8s} " + " ".join(f"r{r}:{m:.4f}" for r, m, _ in rows) + f"   [{p}]")
        if plt and res:
            fig, ax = plt.subplots(figsize=(6, 4))
            for (mt, ut, _), rows in sorted(res.items()):
                ax.plot([r for r, _, _ in rows], [m for _, m, _ in rows], marker="o", label=f"{mt}+{ut}")
            ax.set_xlabel("round"); ax.set_ylabel("mean client AUC"); ax.legend(fontsize=7)
            fig.savefig(a.out or "auc_per_round.pdf", bbox_inches="tight")
    else:
        lat = pickle.load(open(a.path, "rb"))
        rounds = sorted(lat.keys())
        rnd = rounds[a.round] if len(rounds) >= abs(a.round) else rounds[-1]
        devs = sorted(lat[rnd].keys())
        print(f"round {rnd}: {len(devs)} devices; latent dim {lat[rnd][devs[0]][0].shape[1]}")
        if plt:
            try:
                from sklearn.manifold import TSNE
                embed = lambda z: TSNE(n_components=2, init="pca", perplexity=min(30, len(z) - 1)).fit_transform(z)
            except Exception:
                embed = lambda z: (z - z.mean(0)) @ np.linalg.svd(z - z.mean(0), full_matrices=False)[2][:2].T
            n = min(len(devs), 6)
            fig, axes = plt.subplots(1, n, figsize=(3 * n, 3))
            for ax, d in zip(np.atleast_1d(axes), devs[:n]):
                z, y = lat[rnd][d]
                idx = np.random.default_rng(0).permutation(len(z))[:1500]
                e = embed(z[idx])
                ax.scatter(e[:, 0], e[:, 1], c=y[idx], s=2, cmap="coolwarm"); ax.set_title(d, fontsize=7)
            fig.savefig(a.out or "latents.pdf", bbox_inches="tight")


if __name__ == "__main__":
    main()
