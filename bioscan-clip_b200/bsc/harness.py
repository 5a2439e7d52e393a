"""Synthetic full train step around the fused loss (BASELINE config 5).

The caller side of the hot path: ``train_epoch`` (bioscanclip/epoch/train_epoch.py:21-44) driving
``SimpleCLIP.forward`` (bioscanclip/model/simple_clip.py:27-50) with random-init encoders of the reference's shapes
 -- ViT-B/16 (12 x 768, 197 tokens; timm is absent, so the HF ``ViTModel`` of the same geometry), BarcodeBERT
(BERT-base-shaped ``BertForMaskedLM`` over the 5-mer vocabulary with the decoder replaced by ``Linear(768, 768)`` and
the ``softmax.mean(1)`` head, dna_encoder.py:93-105) and BERT-small (4 x 512) + ``Linear(512, 768)``
(language_encoder.py:78-89) -- LoRA r=4 on the attention query / value projections (image_encoder.py:42-48,
dna_encoder.py:40-49), base weights frozen.  Everything here is plain PyTorch: the product is the loss the step calls.
Unlike the reference (which never synchronises gradients across ranks) ``train_step`` all-reduces
the LoRA gradients when a process group is given.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class LoRALinear(nn.Module):
    """y = W x + B(A x) with W frozen (the _LoRALayer of dna_encoder.py / language_encoder.py)."""

    def __init__(self, base: nn.Linear, r: int):
        super().__init__()
        self.base = base
        for p in self.base.parameters():
            p.requires_grad_(False)
        self.a = nn.Linear(base.in_features, r, bias=False)
        self.b = nn.Linear(r, base.out_features, bias=False)
        nn.init.kaiming_uniform_(self.a.weight, a=math.sqrt(5))
        nn.init.zeros_(self.b.weight)

    def forward(self, x):
        return self.base(x) + self.b(self.a(x))


def add_lora_to_attention(module: nn.Module, r: int = 4) -> int:
    """Freeze ``module`` and wrap every attention ``query`` / ``value`` projection; returns the number wrapped."""
    for p in module.parameters():
        p.requires_grad_(False)
    n = 0
    for parent in list(module.modules()):
        for name in ("query", "value"):
            child = getattr(parent, name, None)
            if isinstance(child, nn.Linear):
                setattr(parent, name, LoRALinear(child, r))
                n += 1
    return n


class ImageEncoder(nn.Module):
    def __init__(self, r=4, layers=12):
        super().__init__()
        from transformers import ViTConfig, ViTModel

        self.vit = ViTModel(ViTConfig(hidden_size=768, num_hidden_layers=layers, num_attention_heads=12,
                                      intermediate_size=3072, image_size=224, patch_size=16), add_pooling_layer=False)
        add_lora_to_attention(self.vit, r)

    def forward(self, x):
        return self.vit(pixel_values=x).last_hidden_state[:, 0]  # class token after the final norm


class DnaEncoder(nn.Module):
    VOCAB = 4 ** 5 + 3  # 5-mers over ACGT + special tokens (dna_encoder.py:25-35)

    def __init__(self, r=4, layers=12):
        super().__init__()
        from transformers import BertConfig, BertForMaskedLM

        self.bert = BertForMaskedLM(BertConfig(vocab_size=self.VOCAB, hidden_size=768, num_hidden_layers=layers,
                                               num_attention_heads=12, intermediate_size=3072,
                                               max_position_embeddings=512))
        add_lora_to_attention(self.bert, r)
        self.bert.cls.predictions.decoder = nn.Linear(768, 768)  # trainable head (dna_encoder.py:93-96)

    def forward(self, tokens):
        return self.bert(tokens).logits.softmax(dim=-1).mean(dim=1)  # dna_encoder.py:105


class LanguageEncoder(nn.Module):
    def __init__(self, r=4, layers=4):
        super().__init__()
        from transformers import BertConfig, BertModel

        self.bert = BertModel(BertConfig(vocab_size=30522, hidden_size=512, num_hidden_layers=layers,
                                         num_attention_heads=8, intermediate_size=2048), add_pooling_layer=False)
        add_lora_to_attention(self.bert, r)
        self.proj = nn.Linear(512, 768)

    def forward(self, x):
        return self.proj(self.bert(**x).last_hidden_state.mean(dim=1))  # language_encoder.py:89


class SimpleCLIP(nn.Module):
    """bioscanclip/model/simple_clip.py:18-50 without the open_clip branch."""

    def __init__(self, image_encoder, dna_encoder, language_encoder):
        super().__init__()
        self.image_encoder = image_encoder
        self.dna_encoder = dna_encoder
        self.language_encoder = language_encoder

    def forward(self, image_input, dna_input, language_input):
        image_output = dna_output = language_output = None
        if self.dna_encoder is not None:
            dna_output = F.normalize(self.dna_encoder(dna_input), p=2, dim=-1)
        if self.image_encoder is not None:
            image_output = F.normalize(self.image_encoder(image_input), p=2, dim=-1)
        if self.language_encoder is not None:
            language_output = F.normalize(self.language_encoder(language_input), p=2, dim=-1)
        return image_output, dna_output, language_output


def build_synthetic_model(r=4, vit_layers=12, dna_layers=12, text_layers=4) -> SimpleCLIP:
    return SimpleCLIP(ImageEncoder(r, vit_layers), DnaEncoder(r, dna_layers), LanguageEncoder(r, text_layers))


def synthetic_batch(batch: int, device, seed: int = 0, text_len: int = 20, dna_len: int = 133):
    """rand(B,3,224,224) images, random 660-bp barcodes as 5-mer token ids (660 / 5 = 132 tokens + CLS), random text."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    image = torch.rand(batch, 3, 224, 224, generator=g)
    bases = torch.randint(0, 4, (batch, 660), generator=g)
    kmers = (bases.view(batch, 132, 5) * (4 ** torch.arange(4, -1, -1))).sum(-1) + 3
    dna = torch.cat([torch.zeros(batch, 1, dtype=torch.long), kmers], dim=1)[:, :dna_len]  # id 0 leads (appendix A.10)
    ids = torch.randint(1000, 30000, (batch, text_len), generator=g)
    text = {"input_ids": ids, "token_type_ids": torch.zeros_like(ids), "attention_mask": torch.ones_like(ids)}
    labels = torch.arange(batch)
    return (image.to(device), dna.to(device), {k: v.to(device) for k, v in text.items()}, labels.to(device))


def train_step(model, criterion, optimizer, batch, autocast_dtype=torch.bfloat16, group=None, loss_args=()):
    """One iteration of train_epoch (train_epoch.py:29-40): zero_grad, forward, loss, backward, step."""
    image, dna, text, labels = batch
    optimizer.zero_grad(set_to_none=True)
    with torch.autocast("cuda", dtype=autocast_dtype, enabled=autocast_dtype is not None):
        image_out, dna_out, text_out = model(image, dna, text)
    loss = criterion(image_out, dna_out, text_out, labels, *loss_args)
    loss.backward()
    if group is not None:
        import torch.distributed as dist

        world = dist.get_world_size(group)
        for p in model.parameters():
            if p.requires_grad and p.grad is not None:
                dist.all_reduce(p.grad, group=group)
                p.grad /= world
    optimizer.step()
    return loss.detach()
