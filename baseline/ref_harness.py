"""Reference arm of the benchmark: the UNMODIFIED reference (``baseline/_ref``) driven through its
own public API on the same synthetic scenario, with NCCL for the peer exchange when N > 1.

What runs here is the reference's stock code path and nothing of this framework's engine, models or
kernels: ``DataLoader.IoTDataProccessor`` / ``IoTDataset`` + ``torch.utils.data.DataLoader`` (batch 12,
pinned, per-batch ``.to(device)``), ``Model.Shrink_Autoencoder``, ``Trainer.ClientTrainer.run`` /
``vote_for_aggregator`` / ``aggregate_models`` / ``receive_model`` / ``update_from_peers`` and
``Evaluator.Evaluator.evaluate``, called in the order of the reference's round loop
(``src/main.py:267-340``).  Only the synthetic traffic generator (there is no dataset on the box) and
the wall-clock bracket are ours.

Multi-GPU ("the reference's NCCL build", BASELINE.json): the reference has no distributed code, so the
harness does what a minimal port would: clients are block-partitioned over ranks; each rank trains and
evaluates only the clients it owns; after local training the selected clients' state_dicts are
exchanged with one ``dist.all_gather`` (NCCL) and loaded into each rank's replicas, after which every
rank runs the reference's vote / aggregate / broadcast / verify code unchanged (replicated, as every
peer would); per-client AUCs are all-gathered.
"""
from __future__ import annotations

import os
import random
import shutil
import sys
import tempfile
import time
from typing import Dict, List

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


REF_SRC = os.environ.get("FEDMSE_REFERENCE_SRC", "/root/reference")
_REF_PKGS = ("Model", "Trainer", "Evaluator", "DataLoader", "Utils")


def client_seed(seed0: int, i: int) -> int:
    """numpy seed the harness sets before client ``i``'s shuffles (same on every rank)."""
    return (int(seed0) * 1000003 + i) % (2 ** 32)


def dev_seed(seed0: int) -> int:
    """numpy seed the harness sets before the dev-set sampling."""
    return (int(seed0) * 1000003 + 999983) % (2 ** 32)


def _installed() -> bool:
    return all(os.path.isfile(os.path.join(REF_DIR, p, "__init__.py")) for p in _REF_PKGS)


def ensure_installed(verbose: bool = False) -> str:
    """Idempotent offline install of the unmodified reference into ``baseline/_ref``.

    Order: (1) already there -> nothing to do; (2) ``baseline/install_reference.sh`` (the pip route recorded in
    DESIGN.md section 5); (3) if pip fails, a verbatim file copy of the reference's python packages.  Returns ''
    on success, else the reason.  ``__graft_entry__.build()`` and ``bench.py --impl reference`` both call this, so a
    fresh checkout (``baseline/_ref`` is git-ignored) gets the reference arm without a manual step."""
    if _installed():
        return ""
    src = os.path.join(REF_SRC, "src")
    if not os.path.isdir(os.path.join(src, "Trainer")):
        return f"baseline/_ref missing and no reference source at {REF_SRC}"
    import subprocess
    why = ""
    try:
        r = subprocess.run(["bash", os.path.join(HERE, "install_reference.sh"), REF_SRC], capture_output=True,
                           text=True, timeout=600)
        if r.returncode != 0:
            why = (r.stderr or r.stdout).strip().splitlines()[-1:] or ["pip install failed"]
            why = str(why[0])[:200]
        if verbose:
            print(r.stdout[-2000:], r.stderr[-2000:])
    except Exception as e:                                   # no bash / pip / timeout
        why = repr(e)[:200]
    if not _installed():                                     # fallback: plain copy of the same files
        try:
            os.makedirs(REF_DIR, exist_ok=True)
            for pkg in _REF_PKGS:
                dst = os.path.join(REF_DIR, pkg)
                os.makedirs(dst, exist_ok=True)
                for f in os.listdir(os.path.join(src, pkg)):
                    if f.endswith(".py") and f != "spliter.py":      # spliter.py is a C comment block, not python
                        shutil.copyfile(os.path.join(src, pkg, f), os.path.join(dst, f))
            shutil.copyfile(os.path.join(src, "main.py"), os.path.join(REF_DIR, "fedmse_reference_main.py"))
            with open(os.path.join(REF_DIR, "INSTALL_NOTE.txt"), "w") as fh:
                fh.write(f"plain copy of {src} (pip route failed: {why})\n")
        except Exception as e:
            return f"baseline/_ref install failed: {why}; copy failed: {e!r}"
    return "" if _installed() else f"baseline/_ref install failed: {why}"


def available() -> str:
    """'' when the reference install is usable (installing it first if needed), else the reason."""
    if int(os.environ.get("LOCAL_RANK", "0")) != 0 and not _installed():
        # torchrun: local rank 0 installs, the others wait for it (or for its verdict that it cannot)
        src_ok = os.path.isdir(os.path.join(REF_SRC, "src", "Trainer"))
        t0 = time.time()
        while src_ok and not _installed() and time.time() - t0 < 180:
            time.sleep(1.0)
        time.sleep(1.0)
        return "" if _installed() else "baseline/_ref missing (local rank 0 could not install it)"
    return ensure_installed()


def _import_ref():
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    import logging
    logging.disable(logging.CRITICAL)            # the reference logs every epoch through the root logger
    import DataLoader as RDL
    import Evaluator as REV
    import Model as RM
    import Trainer as RT
    import tqdm
    import Evaluator.evaluator as _ev
    _ev.tqdm = lambda it, **kw: it               # silence the progress bars only (no behavioural change)
    return RDL, REV, RM, RT


class ReferenceFederation:
    def __init__(self, raw_clients, owned: range, cfg: Dict, world: int, rank: int, device):
        from torch.utils.data import ConcatDataset, DataLoader
        RDL, REV, RM, RT = _import_ref()
        self.REV, self.cfg, self.world, self.rank, self.owned, self.device = REV, cfg, world, rank, owned, device
        self.tmp = tempfile.mkdtemp(prefix="fedmse_ref_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
        cwd = os.getcwd()
        os.chdir(self.tmp)                       # the reference writes Checkpoint/ relative to the CWD
        self.client_info, self.trainers = [], []
        import pandas as pd
        dev_parts = []
        bs = cfg["batch_size"]
        seed0 = int(cfg.get("seed", 1234))
        for i, cl in enumerate(raw_clients):     # src/main.py:131-207
            # numpy's global generator drives DataFrame.sample as in the reference; it is re-seeded per client so
            # that the shuffles of a client are the same on every rank (ranks only shuffle the abnormal rows of the
            # clients they own, which would otherwise shift the stream of all later clients and give every rank a
            # different replica of the validation / dev data)
            np.random.seed(client_seed(seed0, i))
            normal = pd.DataFrame(cl.normal).sample(frac=1).reset_index(drop=True)
            n = len(normal)
            n_tr, n_va, n_dv = int(0.4 * n), int(0.1 * n), int(0.4 * n)
            tr, va = normal[:n_tr], normal[n_tr:n_tr + n_va]
            dv, te = normal[n_tr + n_va:n_tr + n_va + n_dv], normal[n_tr + n_va + n_dv:]
            proc = RDL.IoTDataProccessor(scaler="standard")
            p_tr, l_tr = proc.fit_transform(tr)
            p_va, l_va = proc.transform(va)
            info = {"device": cl.name, "dev_normal_dataset": dv, "validation_data": torch.Tensor(p_va),
                    "n_train": len(p_tr), "n_valid": len(p_va)}
            if i in owned:
                abnormal = pd.DataFrame(cl.abnormal).sample(frac=1).reset_index(drop=True)
                p_te, l_te = proc.transform(te)
                p_ab, l_ab = proc.transform(abnormal, type="abnormal")
                p_new, l_new = proc.transform(pd.DataFrame(cl.test_normal))
                p_te = np.concatenate([p_te, p_new], axis=0)
                l_te = np.concatenate([l_te, l_new], axis=0)
                test_ds = ConcatDataset([RDL.IoTDataset(p_te, l_te), RDL.IoTDataset(p_ab, l_ab)])
                info["train_loader"] = DataLoader(RDL.IoTDataset(p_tr, l_tr), batch_size=bs,
                                                  pin_memory=torch.cuda.is_available())
                info["valid_loader"] = DataLoader(RDL.IoTDataset(p_va, l_va), batch_size=bs, pin_memory=True)
                info["test_loader"] = DataLoader(test_ds, batch_size=bs, pin_memory=True)
            self.client_info.append(info)
        np.random.seed(dev_seed(seed0))
        min_len = min(len(c["dev_normal_dataset"]) for c in self.client_info)     # src/main.py:214-223
        dev = pd.concat([c["dev_normal_dataset"].sample(n=min_len) for c in self.client_info], axis=0)
        processed_dev, _ = RDL.IoTDataProccessor(scaler="standard").fit_transform(dev)
        for i, info in enumerate(self.client_info):                                # src/main.py:225-257
            if cfg["model_type"] == "hybrid":
                model = RM.Shrink_Autoencoder(input_dim=115, output_dim=115, shrink_lambda=cfg["shrink_lambda"])
            else:
                model = RM.Autoencoder(input_dim=115, output_dim=115)
            tr = RT.ClientTrainer(model=model, loss_function=torch.nn.MSELoss, optimizer=torch.optim.Adam,
                                  epoch=cfg["epoch"], batch_size=bs, lr_rate=cfg["lr_rate"],
                                  update_type=cfg["update_type"], patience=cfg["patience"],
                                  save_dir=os.path.join(self.tmp, "ClientModel", info["device"]), client_id=i,
                                  model_type=cfg["model_type"], verification_method="val",
                                  verification_threshold=3.0, performance_threshold=0.002)
            tr.create_dev_dataset({"dataset": processed_dev})
            self.trainers.append(tr)
        last_valid = self.client_info[-1]["validation_data"]
        for i, tr in enumerate(self.trainers):                                     # src/main.py:260-264
            tr.connect_to_peers([c for j, c in enumerate(self.trainers) if j != i])
            tr.validation_data = last_valid
        os.chdir(cwd)
        self.samples = 0

    def close(self):
        shutil.rmtree(self.tmp, ignore_errors=True)

    # ---------------------------------------------------------------- one round (src/main.py:267-340)
    def run_round(self, rnd: int) -> List[float]:
        import pickle
        import torch.distributed as dist
        cfg, trainers, info = self.cfg, self.trainers, self.client_info
        n_sel = max(1, int(cfg["num_participants"] * len(trainers)))
        sel_idx = random.sample(range(len(trainers)), n_sel)
        sel_tr = [trainers[i] for i in sel_idx]
        sel_info = [info[i] for i in sel_idx]
        for k, i in enumerate(sel_idx):
            if i in self.owned:
                sel_tr[k].run(sel_info[k]["train_loader"], sel_info[k]["valid_loader"])
                with open(os.path.join(sel_tr[k].save_dir, "training_tracking.pkl"), "rb") as fh:
                    self.samples += len(pickle.load(fh)) * (sel_info[k]["n_train"] + sel_info[k]["n_valid"])
        if self.world > 1:   # peer exchange: the selected clients' models, one NCCL all_gather
            flat = torch.zeros(n_sel, 6764, device=self.device)
            for k, i in enumerate(sel_idx):
                if i in self.owned:
                    flat[k] = torch.cat([p.detach().reshape(-1) for p in sel_tr[k].model.state_dict().values()])
            parts = [torch.empty_like(flat) for _ in range(self.world)]
            dist.all_gather(parts, flat)
            full = torch.stack(parts).sum(0)
            for k, i in enumerate(sel_idx):
                if i not in self.owned:
                    sd, o = sel_tr[k].model.state_dict(), 0
                    for key, t in sd.items():
                        t.copy_(full[k, o:o + t.numel()].reshape(t.shape))
                        o += t.numel()
        aggregator = None
        for client in sel_tr:
            cand = client.vote_for_aggregator(sel_tr, sel_info[0]["validation_data"], rnd)
            if cand:
                aggregator = cand
                break
        if aggregator:
            state = aggregator.aggregate_models(sel_tr, sel_info[0]["validation_data"], rnd)
            if state:
                # src/main.py:298-306 hands the state to every client.  A rank only ever reads the models of the clients
                # it owns (their next training, their evaluation) and of the SELECTED clients (refreshed from their
                # owners by the all_gather above), so with world > 1 it runs the reference's receive + CPU verification
                # for its own clients only -- not N times the same 1000 verifications on every rank.
                mine = trainers if self.world == 1 else [trainers[i] for i in self.owned]
                for client in mine:
                    if client != aggregator:
                        client.receive_model(aggregator, state)
                for client in mine:
                    if client != aggregator:
                        client.update_from_peers()
        metrics = []
        for i in self.owned:
            ev = self.REV.Evaluator(trainers[i].model, model_type=cfg["model_type"], metric="AUC")
            m = ev.evaluate(info[i]["test_loader"], info[i]["train_loader"])
            metrics.append(float(m[0] if isinstance(m, tuple) else m))
        if self.world > 1:
            t = torch.tensor(metrics, device=self.device, dtype=torch.float64)
            parts = [torch.empty_like(t) for _ in range(self.world)]
            dist.all_gather(parts, t)
            metrics = torch.cat(parts).tolist()
        return metrics
