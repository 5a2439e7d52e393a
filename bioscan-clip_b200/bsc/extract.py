"""Embedding extraction with a device-resident bank.

Same names, arguments and return shapes as the reference:
  * ``get_feature_and_label(dataloader, model, device, type_of_feature="dna", for_open_clip=False, multi_gpu=False)``
    -> ``(file_name_list, encoded_feature, label_list)``        bioscanclip/epoch/inference_epoch.py:8-68
  * ``get_features_and_label(dataloader, model, device, for_key_set=False, for_open_clip=False)`` -> dict with the
    reference's keys                                            scripts/inference_and_eval.py:734-783

The reference moves every batch to the host as Python lists (``.cpu().tolist()``, inference_epoch.py:45,57,61) and
builds a float64 array at the end (:67).  Here the encoder output is L2-normalised by ``bsc_l2norm_fwd`` straight
into a preallocated float32 bank in HBM; ``device_resident=True`` returns that bank (ready for
``bsc.IndexFlatIP.add`` without a host round trip), the default returns the reference's float64 numpy array with
the same values.  The encoders themselves are the caller's modules: this file is the caller side of the hot path.
"""
from __future__ import annotations

import numpy as np
import torch

from ._lib import BscError


def convert_label_dict_to_list_of_dict(label_batch):
    """bioscanclip/epoch/eval_epoch.py:26-38."""
    return [{"order": o, "family": f, "genus": g, "species": s}
            for o, f, g, s in zip(label_batch["order"], label_batch["family"], label_batch["genus"],
                                  label_batch["species"])]


class _Bank:
    """float32 [capacity, d] on the device, grown geometrically; rows are written in place."""

    def __init__(self, device, capacity_hint: int):
        self.device = device
        self.hint = max(int(capacity_hint), 1)
        self.buf = None
        self.n = 0

    def slot(self, rows: int, d: int) -> torch.Tensor:
        if self.buf is None:
            self.buf = torch.empty((max(self.hint, rows), d), dtype=torch.float32, device=self.device)
        elif self.n + rows > self.buf.shape[0]:
            grown = torch.empty((max(2 * self.buf.shape[0], self.n + rows), d), dtype=torch.float32,
                                device=self.device)
            grown[:self.n] = self.buf[:self.n]
            self.buf = grown
        out = self.buf[self.n:self.n + rows]
        self.n += rows
        return out

    def result(self):
        return None if self.buf is None else self.buf[:self.n]


def get_feature_and_label(dataloader, model, device, type_of_feature="dna", for_open_clip=False, multi_gpu=False,
                          device_resident=False, ops=None):
    if type_of_feature not in ["dna", "image", "text"]:
        raise TypeError(f"{type_of_feature} is not a valid input type")
    if type_of_feature == "image" and model.image_encoder is None:
        return None, None, None
    if type_of_feature == "dna" and model.dna_encoder is None:
        return None, None, None
    if type_of_feature == "text" and model.language_encoder is None:
        return None, None, None
    device = torch.device(device)
    if device.type != "cuda":
        raise BscError("bsc.get_feature_and_label writes the bank on a CUDA device; there is no CPU path")
    if ops is None:
        from .ops import CudaOps

        ops = CudaOps()
    try:
        hint = len(dataloader.dataset)
    except Exception:
        hint = 1024
    bank = _Bank(device, hint)
    label_list, file_name_list = [], []
    model.eval()
    with torch.no_grad():
        for batch in dataloader:
            processid_batch, image_input_batch, dna_input_batch, input_ids, token_type_ids, attention_mask, \
                label_batch = batch
            if type_of_feature == "text":
                if for_open_clip:
                    language_input = input_ids
                else:
                    language_input = {"input_ids": input_ids.to(device), "token_type_ids": token_type_ids.to(device),
                                      "attention_mask": attention_mask.to(device)}
                raw, normalise = model.language_encoder(language_input), True
            else:
                x = (dna_input_batch if type_of_feature == "dna" else image_input_batch).to(device)
                enc_owner = model.module if multi_gpu else model
                enc = enc_owner.dna_encoder if type_of_feature == "dna" else enc_owner.image_encoder
                # the multi_gpu branch of the reference does not normalise (inference_epoch.py:38-39, :50-51)
                raw, normalise = (x if enc is None else enc(x)), not multi_gpu
            raw = raw.reshape(raw.shape[0], -1)
            slot = bank.slot(raw.shape[0], raw.shape[1])
            if normalise:
                ops.normalize(raw if raw.dtype in (torch.float32, torch.bfloat16, torch.float16) else raw.float(),
                              torch.float32, out=slot)
            else:
                slot.copy_(raw)
            label_list = label_list + convert_label_dict_to_list_of_dict(label_batch)
            file_name_list = file_name_list + list(processid_batch)
    feats = bank.result()
    if feats is None:
        return file_name_list, (None if device_resident else np.array([])), label_list
    if device_resident:
        return file_name_list, feats, label_list
    return file_name_list, feats.cpu().numpy().astype(np.float64), label_list


def get_features_and_label(dataloader, model, device, for_key_set=False, for_open_clip=False, device_resident=False):
    """scripts/inference_and_eval.py:734-783 with the reference's dictionary keys; with ``device_resident=True`` the
    feature entries are device tensors and mean / concatenation run on the device."""
    model.eval()
    kw = dict(multi_gpu=False, for_open_clip=for_open_clip, device_resident=device_resident)
    _, lang, _ = get_feature_and_label(dataloader, model, device, type_of_feature="text", **kw)
    _, dna, _ = get_feature_and_label(dataloader, model, device, type_of_feature="dna", **kw)
    file_name_list, img, label_list = get_feature_and_label(dataloader, model, device, type_of_feature="image", **kw)
    averaged = concatenated = all_key_features = all_key_features_label = None
    if dna is not None and img is not None:
        if device_resident:
            averaged = (img + dna) / 2
            concatenated = torch.cat((img, dna), dim=1)
        else:
            averaged = np.mean([img, dna], axis=0)
            concatenated = np.concatenate((img, dna), axis=1)
    out = {"file_name_list": file_name_list, "encoded_dna_feature": dna, "encoded_image_feature": img,
           "encoded_language_feature": lang, "averaged_feature": averaged, "concatenated_feature": concatenated,
           "label_list": label_list}
    if for_key_set and img is not None and dna is not None and lang is not None:
        parts = [img, dna, lang]
        all_key_features = torch.cat(parts, dim=0) if device_resident else np.concatenate(parts, axis=0)
        all_key_features_label = label_list + label_list + label_list
    out["all_key_features"] = all_key_features
    out["all_key_features_label"] = all_key_features_label
    return out
