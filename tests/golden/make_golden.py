"""Generate golden fixtures by EXECUTING the reference in the build container.

Run once here (``/root/reference`` is read-only and absent on the GPU box); the .npz files it writes are
committed and are what the tests read.

    python tests/golden/make_golden.py

Loss: imports ``bioscanclip.model.loss_func`` from /root/reference (needs torch only) and records
ContrastiveLoss forward + autograd gradients in float32 exactly as scripts/train_cl.py:190 builds it.
ClipLoss: runs the reference under a 2-process gloo group and records every rank's loss and local grads.
Retrieval: faiss is not installable, so only ``sklearn.preprocessing.normalize`` (inference_and_eval.py:416)
is recorded.
"""
import os
import sys

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def synth(n, d, m_present, seed, normalised, labels_kind):
    g = torch.Generator().manual_seed(seed)
    z = torch.randn(n, d, generator=g)
    feats = []
    for m in range(3):
        if not m_present[m]:
            feats.append(None)
            continue
        x = z + 0.5 * torch.randn(n, d, generator=g)
        if normalised:
            x = torch.nn.functional.normalize(x, dim=1)
        else:
            x = x * (0.5 + 1.5 * torch.rand(n, 1, generator=g))
        feats.append(x)
    if labels_kind == "arange":
        labels = torch.arange(n)
    else:
        labels = torch.randint(0, max(2, n // 4), (n,), generator=g)
    return feats, labels


CASES = {
    # name: (n, d, present, seed, normalised, labels)
    "pair_small": (32, 16, (True, True, False), 1, True, "arange"),
    "tri_multi": (48, 64, (True, True, True), 2, False, "multi"),
    "tri_768": (64, 768, (True, True, True), 3, True, "multi"),
    "img_txt_768": (40, 768, (True, False, True), 4, False, "arange"),
    "ragged_tri": (37, 24, (True, True, True), 5, False, "multi"),
}


def run_loss_cases():
    sys.path.insert(0, REF)
    from bioscanclip.model.loss_func import ContrastiveLoss  # the reference itself

    for name, (n, d, present, seed, normalised, kind) in CASES.items():
        feats, labels = synth(n, d, present, seed, normalised, kind)
        leaves = [None if f is None else f.clone().requires_grad_(True) for f in feats]
        crit = ContrastiveLoss(criterion=nn.CrossEntropyLoss(), logit_scale=1 / 0.07)
        loss = crit(leaves[0], leaves[1], leaves[2], labels)
        loss.backward()
        out = {"labels": labels.numpy(), "loss": loss.detach().numpy(), "logit_scale": np.float64(1 / 0.07)}
        for k, nm in enumerate(["image", "dna", "text"]):
            if feats[k] is not None:
                out[f"x_{nm}"] = feats[k].numpy()
                out[f"g_{nm}"] = leaves[k].grad.numpy()
        np.savez_compressed(os.path.join(HERE, f"loss_{name}.npz"), **out)
        print(name, float(loss))


def _clip_worker(rank, world, port, feats, labels, ret):
    import torch.distributed as dist

    sys.path.insert(0, REF)
    from bioscanclip.model.loss_func import ClipLoss

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = labels.shape[0] // world
    sl = slice(rank * n, (rank + 1) * n)
    leaves = [None if f is None else f[sl].clone().requires_grad_(True) for f in feats]
    crit = ClipLoss(rank=rank, world_size=world, criterion=nn.CrossEntropyLoss())
    loss = crit(leaves[0], leaves[1], leaves[2], labels[sl], 1 / 0.07)
    loss.backward()
    ret[rank] = (loss.detach().numpy(), [None if l is None else l.grad.numpy() for l in leaves])
    dist.destroy_process_group()


def run_clip_case():
    import torch.multiprocessing as mp

    n, d = 32, 48
    feats, labels = synth(n, d, (True, True, True), 7, False, "multi")
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_clip_worker, args=(2, 29611, feats, labels, ret), nprocs=2, join=True)
    out = {"labels": labels.numpy(), "logit_scale": np.float64(1 / 0.07)}
    for k, nm in enumerate(["image", "dna", "text"]):
        out[f"x_{nm}"] = feats[k].numpy()
    for r in range(2):
        out[f"loss_rank{r}"] = ret[r][0]
        for k, nm in enumerate(["image", "dna", "text"]):
            out[f"g_{nm}_rank{r}"] = ret[r][1][k]
    np.savez_compressed(os.path.join(HERE, "cliploss_w2.npz"), **out)
    print("cliploss", [float(ret[r][0]) for r in range(2)])


def run_normalize_case():
    from sklearn.preprocessing import normalize

    rng = np.random.default_rng(11)
    x = rng.standard_normal((24, 20))
    x[3] = 0.0
    x[17] *= 1e-30
    y = normalize(x, norm="l2", axis=1).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "retrieval_normalize.npz"), x=x, y=y)


def run_accuracy_case():
    """scripts/inference_and_eval.py cannot be imported (faiss, hydra, h5py ... are missing), so the SOURCE of
    top_k_micro_accuracy / top_k_macro_accuracy and of the label loop of make_prediction is cut out of the file
    with ``ast`` and executed verbatim; inputs and outputs go to accuracy_small.json."""
    import ast
    import json

    path = os.path.join(REF, "scripts", "inference_and_eval.py")
    src = open(path).read()
    tree = ast.parse(src)
    env = {"LEVELS": ["order", "family", "genus", "species"]}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in ("top_k_micro_accuracy", "top_k_macro_accuracy"):
            exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), env)
    rng = np.random.default_rng(5)
    nb, nq, k = 60, 40, 5
    names = {"order": 3, "family": 5, "genus": 9, "species": 14}

    def rand_label():
        return {lv: f"{lv}_{int(rng.integers(0, c))}" for lv, c in names.items()}

    keys_label = [rand_label() for _ in range(nb)]
    gt_list = [rand_label() for _ in range(nq)]
    indices = np.stack([rng.permutation(nb)[:k] for _ in range(nq)]).astype(np.int64)
    # the label loop of make_prediction (inference_and_eval.py:422-433), executed from its own source lines
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "make_prediction")
    loop = next(n for n in fn.body if isinstance(n, ast.For))
    loc = {"indices": indices, "keys_label": keys_label, "pred_list": [], "LEVELS": env["LEVELS"]}
    exec(compile(ast.Module(body=[loop], type_ignores=[]), path, "exec"), loc)
    pred_list = loc["pred_list"]
    k_list = [1, 3, 5]
    micro = env["top_k_micro_accuracy"](pred_list, gt_list, k_list=k_list)
    macro, per_class = env["top_k_macro_accuracy"](pred_list, gt_list, k_list=k_list)
    out = {"keys_label": keys_label, "gt_list": gt_list, "indices": indices.tolist(), "k_list": k_list,
           "pred_list": pred_list,
           "micro": {str(a): b for a, b in micro.items()}, "macro": {str(a): b for a, b in macro.items()},
           "per_class": {str(a): b for a, b in per_class.items()}}
    with open(os.path.join(HERE, "accuracy_small.json"), "w") as f:
        json.dump(out, f)
    print("accuracy_small.json", micro[1], macro[5])


if __name__ == "__main__":
    torch.manual_seed(0)
    torch.set_num_threads(4)
    run_loss_cases()
    run_clip_case()
    run_normalize_case()
    run_accuracy_case()
