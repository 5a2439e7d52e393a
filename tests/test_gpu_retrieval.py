"""GPU parity: bank scan top-k (through the C ABI) vs the retrieval oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import retrieval_oracle as ro
from tests.synth import bank_inputs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nb,nq,d,k", [(5000, 257, 768, 5), (70000, 64, 128, 5), (333, 10, 40, 3), (3, 4, 16, 5)])
def test_fp32_indices_bit_exact(nb, nq, d, k):
    import bsc

    bank, q, pi = bank_inputs(nb, nq, d, seed=nb)
    ref_s, ref_i = ro.search_ip(bank, q, k)
    index = bsc.IndexFlatIP(d, precision="fp32")
    index.add(bank)
    s, i = index.search(q, k)
    assert s.dtype == np.float32 and i.dtype == np.int64 and i.shape == (nq, k)
    np.testing.assert_array_equal(i, ref_i)
    np.testing.assert_allclose(s, ref_s, rtol=0, atol=3e-6)


@pytest.mark.parametrize("nb,nq,d,k", [(4096, 128, 768, 5), (10000, 300, 768, 5), (777, 33, 128, 1), (70000, 32, 768, 5)])
def test_bf16_scan_matches_oracle_on_rounded_bank(nb, nq, d, k):
    """bf16 mode: the oracle sees the same bf16-rounded bank and queries, so planted neighbours and exact
    duplicates must come back with identical ids; scores agree to fp32 accumulation error."""
    import bsc

    bank, q, pi = bank_inputs(nb, nq, d, seed=nb + 1)
    bank_r = torch.from_numpy(bank).bfloat16().float().numpy()
    q_r = torch.from_numpy(q).bfloat16().float().numpy()
    ref_s, ref_i = ro.search_ip(bank_r, q_r, k)
    index = bsc.IndexFlatIP(d, precision="bf16")
    index.add(bank)
    s, i = index.search(q, k)
    # top-1 is planted (margin >> rounding); deeper ranks may swap only inside fp32 near-ties
    np.testing.assert_array_equal(i[:, 0], ref_i[:, 0])
    np.testing.assert_allclose(s, ref_s, rtol=0, atol=2e-5)
    mism = i != ref_i
    if mism.any():
        gap = np.abs(s - ref_s)[mism]
        assert gap.max() <= 2e-5


@pytest.mark.parametrize("parts,nq,k", [(444, 32, 5), (888, 7, 8), (60, 40, 5), (3, 5000, 5), (500, 3, 1), (51, 2048, 5)])
def test_topk_merge_matches_numpy(parts, nq, k):
    """bsc_topk_merge (block-per-query kernel for few queries and many lists, warp-per-query kernel otherwise) against a
    numpy sort on (score desc, id asc): scores drawn from a handful of values so that ties are everywhere, a third of
    the entries padding (id -1), one query with fewer than k valid candidates."""
    from bsc.ops import CudaOps

    rng = np.random.default_rng(parts * 1000 + nq)
    lowest = np.float32(-3.4028234663852886e38)
    total = parts * k
    scores = rng.integers(0, 12, size=(nq, total)).astype(np.float32) / 4
    ids = np.stack([rng.permutation(total * 3)[:total] for _ in range(nq)]).astype(np.int64)
    pad = rng.random((nq, total)) < 0.33
    pad[0] = True
    pad[0, :max(1, k - 2)] = False  # query 0: fewer than k candidates
    ids[pad] = -1
    scores[pad] = lowest
    want_s = np.full((nq, k), lowest, dtype=np.float32)
    want_i = np.full((nq, k), -1, dtype=np.int64)
    for q in range(nq):
        ok = np.nonzero(ids[q] >= 0)[0]
        order = ok[np.lexsort((ids[q][ok], -scores[q][ok]))][:k]
        want_s[q, :len(order)] = scores[q][order]
        want_i[q, :len(order)] = ids[q][order]
    cs = torch.from_numpy(scores.reshape(nq, parts, k).transpose(1, 0, 2).copy()).cuda()
    ci = torch.from_numpy(ids.reshape(nq, parts, k).transpose(1, 0, 2).copy()).cuda()
    got_s, got_i = CudaOps().topk_merge(cs, ci)
    np.testing.assert_array_equal(got_i.cpu().numpy(), want_i)
    np.testing.assert_array_equal(got_s.cpu().numpy(), want_s)


def test_make_prediction_forms_and_labels():
    import bsc

    bank, q, pi = bank_inputs(600, 20, 64, seed=4)
    labels = [{lv: f"{lv}_{i % 7}" for lv in ro.LEVELS} for i in range(600)]
    ref = ro.make_prediction(q * 3.0, bank * 0.5, labels, True, True, 5)
    out = bsc.make_prediction(q * 3.0, bank * 0.5, labels, True, True, 5)
    assert out[0] == ref[0]
    np.testing.assert_array_equal(out[2], ref[2])
    np.testing.assert_allclose(out[1], ref[1], atol=3e-6)
    only = bsc.make_prediction(q, bank, labels, max_k=3)
    assert isinstance(only, list) and len(only) == 20 and len(only[0]["species"]) == 3


def test_merge_across_incremental_adds_and_ties():
    import bsc

    bank, q, pi = bank_inputs(3000, 50, 96, seed=8)
    ref_s, ref_i = ro.search_ip(bank, q, 5)
    index = bsc.IndexFlatIP(96)
    index.add(bank[:1000])
    index.add(bank[1000:])
    assert index.ntotal == 3000
    s, i = index.search(q, 5)
    np.testing.assert_array_equal(i, ref_i)


def test_full_size_properties():
    """625k x 768 bf16 shard (one GPU's share of the 5M bank): every query must find its planted row first,
    and searching the two halves separately then merging must equal the one-pass search."""
    import bsc

    nb, nq, d = 625_000, 256, 768
    g = torch.Generator(device="cuda").manual_seed(3)
    bank = torch.nn.functional.normalize(torch.randn(nb, d, device="cuda", generator=g), dim=1)
    pi = torch.randint(0, nb, (nq,), device="cuda", generator=g)
    q = torch.nn.functional.normalize(bank[pi] + 0.3 / d ** 0.5 * torch.randn(nq, d, device="cuda", generator=g), dim=1)
    index = bsc.IndexFlatIP(d, precision="bf16")
    index.add(bank)
    s, i = index.search_device(q, 5)
    assert (i[:, 0] == pi).all()
    assert (s[:, :-1] >= s[:, 1:]).all()
    two = bsc.IndexFlatIP(d, precision="bf16")
    two.add(bank[: nb // 2])
    two.add(bank[nb // 2:])
    s2, i2 = two.search_device(q, 5)
    assert (i2 == i).all() and torch.equal(s2, s)


def _rand_taxonomy(rng, n, sizes=(4, 9, 30, 80)):
    levels = ["order", "family", "genus", "species"]
    return [{lv: f"{lv}{int(rng.integers(0, c))}" for lv, c in zip(levels, sizes)} for _ in range(n)]


def test_taxonomy_accuracy_matches_reference_fixture():
    """bsc.TaxonomyEval / top_k_*_accuracy vs outputs of the reference's own function source, exact (==)."""
    import json

    import bsc

    z = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "accuracy_small.json")))
    micro, macro, per_class = bsc.TaxonomyEval(z["keys_label"]).evaluate(np.array(z["indices"]), z["gt_list"], z["k_list"])
    micro2 = bsc.top_k_micro_accuracy(z["pred_list"], z["gt_list"], z["k_list"])
    macro2, per2 = bsc.top_k_macro_accuracy(z["pred_list"], z["gt_list"], z["k_list"])
    for k in z["k_list"]:
        assert micro[k] == z["micro"][str(k)] == micro2[k]
        assert macro[k] == z["macro"][str(k)] == macro2[k]
        assert per_class[k] == z["per_class"][str(k)] == per2[k]


@pytest.mark.parametrize("nq,nb,k", [(1, 7, 5), (3000, 5000, 5), (20000, 300, 8)])
def test_taxonomy_accuracy_matches_oracle(nq, nb, k):
    import bsc
    from oracle import accuracy_oracle as ao

    rng = np.random.default_rng(nq)
    keys, gts = _rand_taxonomy(rng, nb), _rand_taxonomy(rng, nq)
    idx = rng.integers(0, nb, size=(nq, k)).astype(np.int64)
    idx[rng.random((nq, k)) < 0.05] = -1  # faiss padding
    k_list = [1, 3, k]
    pred = [{lv: [keys[i][lv] if i >= 0 else None for i in row] for lv in ao.LEVELS} for row in idx]
    want_micro = ao.top_k_micro_accuracy(pred, gts, k_list)
    want_macro, want_pc = ao.top_k_macro_accuracy(pred, gts, k_list)
    micro, macro, pc, pred_ids = bsc.TaxonomyEval(keys).evaluate(torch.from_numpy(idx).cuda(), gts, k_list,
                                                                 with_pred_ids=True)
    assert micro == want_micro and macro == want_macro and pc == want_pc
    assert pred_ids.shape == (nq, k, 4)
    assert bool((pred_ids[torch.from_numpy(idx).cuda() < 0] == -1).all())


class _ToyModel(torch.nn.Module):
    """Three linear 'encoders' with the attribute names get_feature_and_label reads (inference_epoch.py:12-17)."""

    def __init__(self, d_in=24, d=64):
        super().__init__()
        g = torch.Generator().manual_seed(3)
        self.image_encoder = torch.nn.Linear(d_in, d)
        self.dna_encoder = torch.nn.Linear(d_in, d)
        self._lang = torch.nn.Linear(d_in, d)
        for p in self.parameters():
            p.data = torch.randn(p.shape, generator=g)

    def language_encoder(self, inp):
        return self._lang(inp["input_ids"].float())


def _toy_batches(n_batches=3, bs=5, d_in=24):
    g = torch.Generator().manual_seed(9)
    out = []
    for b in range(n_batches):
        ids = [f"p{b}_{i}" for i in range(bs)]
        lab = {lv: [f"{lv}{(b + i) % 3}" for i in range(bs)] for lv in ["order", "family", "genus", "species"]}
        out.append((ids, torch.randn(bs, d_in, generator=g), torch.randn(bs, d_in, generator=g),
                    torch.randint(0, 9, (bs, d_in), generator=g), torch.zeros(bs, d_in), torch.ones(bs, d_in), lab))
    return out


def test_embedding_extraction_device_bank_matches_reference_steps():
    """inference_epoch.py:8-68 restated inline: F.normalize(encoder(x)) per batch, rows in loader order."""
    import bsc

    model = _ToyModel().cuda()
    batches = _toy_batches()
    for kind, col, enc in (("image", 1, model.image_encoder), ("dna", 2, model.dna_encoder)):
        names, feats, labels = bsc.get_feature_and_label(batches, model, "cuda", type_of_feature=kind)
        with torch.no_grad():
            want = torch.cat([torch.nn.functional.normalize(enc(b[col].cuda()), dim=-1) for b in batches]).cpu()
        assert feats.dtype == np.float64 and feats.shape == (15, 64)
        assert np.abs(feats - want.double().numpy()).max() <= 2e-7
        assert names == [n for b in batches for n in b[0]]
        assert labels[6] == {lv: f"{lv}{(1 + 1) % 3}" for lv in ["order", "family", "genus", "species"]}
        _, dev_feats, _ = bsc.get_feature_and_label(batches, model, "cuda", type_of_feature=kind, device_resident=True)
        assert dev_feats.is_cuda and np.array_equal(dev_feats.cpu().numpy().astype(np.float64), feats)
    d = bsc.get_features_and_label(batches, model, "cuda", for_key_set=True, device_resident=True)
    assert set(d) == {"file_name_list", "encoded_dna_feature", "encoded_image_feature", "encoded_language_feature",
                      "averaged_feature", "concatenated_feature", "label_list", "all_key_features",
                      "all_key_features_label"}
    assert d["concatenated_feature"].shape == (15, 128) and d["all_key_features"].shape == (45, 64)
    assert len(d["all_key_features_label"]) == 45
    with pytest.raises(TypeError):
        bsc.get_feature_and_label(batches, model, "cuda", type_of_feature="audio")


def test_stream_to_index_equals_single_add(tmp_path):
    import bsc
    from bsc import storage

    bank, q, _ = bank_inputs(5000, 40, 64, seed=12)
    np.save(tmp_path / "bank.npy", bank)
    mm = np.load(tmp_path / "bank.npy", mmap_mode="r")
    a = bsc.IndexFlatIP(64)
    a.add(bank)
    b = bsc.IndexFlatIP(64)
    assert storage.stream_to_index(b, mm, chunk_rows=1500) == 5000 and b.ntotal == 5000
    sa, ia = a.search(q, 5)
    sb, ib = b.search(q, 5)
    assert np.array_equal(ia, ib) and np.array_equal(sa, sb)


@pytest.mark.parametrize("nb,nq,d,n_random", [(50000, 96, 768, 32), (20000, 300, 128, 300), (700, 40, 64, 10)])
def test_exact_mode_screen_and_rescore_equals_exhaustive_float32(nb, nq, d, n_random):
    """precision='exact' (bf16 screening + float32 re-scoring + per-query proof) returns the indices of the exhaustive
    float32 search: planted neighbours, pure-random queries (small margins) and duplicated bank rows (ties)."""
    import bsc

    bank, q, _ = bank_inputs(nb, nq, d, seed=nb % 17, n_random=n_random)
    want_s, want_i = ro.search_ip(bank, q, 5)
    idx = bsc.IndexFlatIP(d, precision="exact")
    idx.add(bank)
    s, i = idx.search(q, 5)
    assert np.array_equal(i, want_i)
    assert np.abs(s - want_s).max() <= 2e-6
    assert 0 <= idx.last_rechecked <= nq
    ref = bsc.IndexFlatIP(d, precision="fp32")
    ref.add(bank)
    assert np.array_equal(ref.search(q, 5)[1], i)
    # un-normalised rows of mixed scale: the bound scales with the largest row norm
    scale = np.random.default_rng(1).uniform(0.5, 3.0, size=(nb, 1)).astype(np.float32)
    idx2 = bsc.IndexFlatIP(d, precision="exact")
    idx2.add(bank * scale)
    assert np.array_equal(idx2.search(q, 5)[1], ro.search_ip(bank * scale, q, 5)[1])


def test_concatenated_feature_width_1536_and_per_k_searches():
    """The shapes the reference's eval uses besides 768-d: `concatenated_feature` (image + DNA, 1536-d,
    scripts/inference_and_eval.py:753-754) and one search per k in k_list on ONE index
    (bioscanclip/epoch/eval_epoch.py:199-203): k = 1, 3, 5 must be prefixes of each other."""
    import bsc

    bank, q, pi = bank_inputs(9000, 70, 1536, seed=5, n_random=8)
    want_s, want_i = ro.search_ip(bank, q, 5)
    for precision in ("exact", "fp32"):
        index = bsc.IndexFlatIP(1536, precision=precision)
        index.add(bank)
        got = {}
        for k in (1, 3, 5):
            s, i = index.search(q, k)
            assert s.shape == (70, k) and i.dtype == np.int64
            got[k] = i
            np.testing.assert_array_equal(i, want_i[:, :k])
            np.testing.assert_allclose(s, want_s[:, :k], rtol=0, atol=4e-6)
        assert np.array_equal(got[5][:, :1], got[1]) and np.array_equal(got[5][:, :3], got[3])
    index = bsc.IndexFlatIP(1536, precision="bf16")
    index.add(bank)
    np.testing.assert_array_equal(index.search(q[:60], 5)[1][:, 0], pi[:60])  # planted neighbours (random tail excluded)


def test_default_precision_is_the_exact_float32_search():
    """A drop-in user (`import bsc as faiss`) gets float32 results from the tensor-core path, not the slow checker."""
    import bsc

    bank, q, _ = bank_inputs(30000, 64, 768, seed=6, n_random=16)
    index = bsc.IndexFlatIP(768)
    assert index.precision == "exact"
    index.add(bank)
    s, i = index.search(q, 5)
    np.testing.assert_array_equal(i, ro.search_ip(bank, q, 5)[1])
    # the second search of the same shape reuses the cached workspaces: same answer
    s2, i2 = index.search(q, 5)
    assert np.array_equal(i, i2) and np.array_equal(s, s2)


def test_make_prediction_normalises_float64_in_float64():
    """inference_and_eval.py:416-417: normalize(float64 array).astype(float32).  The operands of the search must be
    the SAME float32 numbers the reference feeds faiss, bit for bit (rounding the rows to float32 before normalising
    changes the last bit of about a third of them)."""
    import bsc
    from bsc.retrieval import _normalized_f32
    from bsc.ops import CudaOps

    rng = np.random.default_rng(3)
    x = rng.standard_normal((500, 768)) * rng.uniform(0.1, 30.0, size=(500, 1))
    x[7] = 0.0  # sklearn leaves zero rows alone
    want = ro.l2_normalize_rows(x).astype(np.float32)
    got = _normalized_f32(CudaOps(), x, torch.device("cuda")).cpu().numpy()
    assert np.array_equal(got[7], np.zeros(768, np.float32))
    mism = (got != want).mean()
    # float64 sqrt / division are correctly rounded on both sides; only the summation order of the norm differs
    assert mism <= 1e-4, mism
    naive = ro.l2_normalize_rows(x.astype(np.float32)).astype(np.float32)
    assert (naive != want).mean() > 0.05  # the order of rounding matters: this is what the float64 path avoids
    labels = [{lv: f"{lv}_{i % 11}" for lv in ro.LEVELS} for i in range(500)]
    q = x[:40] + 0.01 * rng.standard_normal((40, 768))
    ref = ro.make_prediction(q, x, labels, True, True, 5)
    out = bsc.make_prediction(q, x, labels, True, True, 5)
    assert out[0] == ref[0]
    np.testing.assert_array_equal(out[2], ref[2])


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("precision", ["exact", "bf16", "fp32"])
def test_index_sharded_over_two_devices_in_one_process(precision):
    """IndexFlatIP(devices=[cuda:0, cuda:1]): function attributes and SM counts are per device, the shards are scanned
    concurrently, duplicates across the shard boundary keep the lowest id."""
    import bsc

    bank, q, pi = bank_inputs(40000, 200, 768, seed=9, n_random=20)
    bank[30000] = bank[100]  # a tie across the two shards
    q[0] = bank[100]
    index = bsc.IndexFlatIP(768, precision=precision, devices=["cuda:0", "cuda:1"])
    index.add(bank)
    s, i = index.search(q, 5)
    if precision == "bf16":
        rb = torch.from_numpy(bank).bfloat16().float().numpy()
        rq = torch.from_numpy(q).bfloat16().float().numpy()
        want = ro.search_ip(rb, rq, 5)[1]
        np.testing.assert_array_equal(i[:, 0], want[:, 0])
    else:
        np.testing.assert_array_equal(i, ro.search_ip(bank, q, 5)[1])
    assert i[0, 0] == 100 and i[0, 1] == 30000
