"""Safetensors checkpoints for the W4A16 weight loaders.

The reference memory-maps `model.safetensors` -- or the shards named by `model.safetensors.index.json`'s
`weight_map` -- into a VarBuilder (loader.rs:189-198, 408-444) and the quantized weight loaders pull
"<prefix>.qweight" / ".qzeros" / ".scales" / ".g_idx" / ".bias" out of it (weight_loader.rs:253-317, 596-714).
`SafetensorsCheckpoint` is that VarBuilder for this repo: a read-only Mapping name -> torch.Tensor over memory-mapped
files, tensors materialised lazily (one layer's packed words at a time: a 4-bit 8B checkpoint is ~5 GB), usable
directly as the `tensors` argument of AwqWeightLoader / GptqWeightLoader / create_weight_loader_with_params.

The file format (8-byte little-endian header length, JSON header {name: {dtype, shape, data_offsets}}, raw bytes) is
read and written here in plain Python + numpy: there is no `safetensors` package in this image, and the writer exists so
that the tests and bench can build small checkpoints of the reference's layout.
"""
from __future__ import annotations

import hashlib
import json
import mmap
import os
import struct
from collections.abc import Mapping
from typing import Dict, Iterator, Optional

import numpy as np
import torch

from ._lib import W4A16Error

# safetensors dtype tag -> (numpy dtype used to view the bytes, torch dtype handed out)
_DTYPES = {
    "F16": (np.uint16, torch.float16), "BF16": (np.uint16, torch.bfloat16), "F32": (np.float32, torch.float32),
    "I32": (np.int32, torch.int32), "U32": (np.int32, torch.int32),   # packed words: same bits, torch lacks a stable u32
    "I64": (np.int64, torch.int64), "U8": (np.uint8, torch.uint8), "I8": (np.int8, torch.int8),
    "I16": (np.int16, torch.int16), "F64": (np.float64, torch.float64), "BOOL": (np.uint8, torch.bool),
}
_TAG_OF = {torch.float16: "F16", torch.bfloat16: "BF16", torch.float32: "F32", torch.int32: "I32",
           torch.int64: "I64", torch.uint8: "U8", torch.int8: "I8", torch.int16: "I16", torch.float64: "F64",
           torch.bool: "BOOL"}


class _File:
    """One .safetensors file: parsed header + a lazily opened read-only mmap."""

    def __init__(self, path: str):
        self.path = path
        size = os.path.getsize(path)
        with open(path, "rb") as f:
            head = f.read(8)
            if len(head) != 8:
                raise W4A16Error(f"{path}: not a safetensors file (shorter than 8 bytes)")
            (n,) = struct.unpack("<Q", head)
            if n > size - 8 or n > (100 << 20):
                raise W4A16Error(f"{path}: header length {n} does not fit the file ({size} bytes)")
            raw = f.read(n)
        try:
            header = json.loads(raw.decode("utf-8"))
        except (UnicodeDecodeError, json.JSONDecodeError) as e:
            raise W4A16Error(f"{path}: header is not JSON: {e}") from e
        if not isinstance(header, dict):
            raise W4A16Error(f"{path}: header is not a JSON object")
        self.metadata = header.pop("__metadata__", None) or {}
        self.data_start = 8 + n
        self.entries: Dict[str, dict] = {}
        payload = size - self.data_start
        for name, ent in header.items():
            try:
                tag, shape, (b, e) = ent["dtype"], [int(d) for d in ent["shape"]], ent["data_offsets"]
            except (KeyError, TypeError, ValueError) as ex:
                raise W4A16Error(f"{path}: malformed header entry for '{name}'") from ex
            if tag not in _DTYPES:
                raise W4A16Error(f"{path}: tensor '{name}' has unsupported dtype {tag}")
            want = int(np.prod(shape, dtype=np.int64)) * np.dtype(_DTYPES[tag][0]).itemsize
            if not (0 <= b <= e <= payload) or e - b != want:
                raise W4A16Error(f"{path}: tensor '{name}' offsets [{b}, {e}) do not match shape {shape} of {tag}")
            self.entries[name] = {"dtype": tag, "shape": shape, "begin": b, "end": e}
        self.header_hash = hashlib.blake2b(raw, digest_size=8).hexdigest() + f":{size}"
        self._mm: Optional[mmap.mmap] = None

    def tensor(self, name: str) -> torch.Tensor:
        ent = self.entries[name]
        if self._mm is None:
            with open(self.path, "rb") as f:
                self._mm = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ)
        np_dt, t_dt = _DTYPES[ent["dtype"]]
        count = (ent["end"] - ent["begin"]) // np.dtype(np_dt).itemsize
        a = np.frombuffer(self._mm, dtype=np_dt, count=count, offset=self.data_start + ent["begin"])
        # one host copy out of the page cache (the mmap is read-only; torch wants writable memory)
        t = torch.from_numpy(a.copy())
        if t_dt in (torch.float16, torch.bfloat16):
            t = t.view(t_dt)
        elif t_dt is torch.bool:
            t = t.to(torch.bool)
        return t.reshape(ent["shape"])


class SafetensorsCheckpoint(Mapping):
    """name -> tensor over one file, a list of files, or a model directory (`model.safetensors`, else the shards of
    `model.safetensors.index.json`, else every *.safetensors in natural order: loader.rs:418-444)."""

    def __init__(self, source):
        paths = self._resolve(source)
        if not paths:
            raise W4A16Error(f"no safetensors files found in {source}")
        self._files = [_File(p) for p in paths]
        self._where: Dict[str, _File] = {}
        for f in self._files:
            for name in f.entries:
                if name in self._where:
                    raise W4A16Error(f"tensor '{name}' appears in both {self._where[name].path} and {f.path}")
                self._where[name] = f
        if self._index_map is not None:   # the index must agree with what the shards hold
            for name, fname in self._index_map.items():
                f = self._where.get(name)
                if f is None or os.path.basename(f.path) != fname:
                    raise W4A16Error(f"index maps '{name}' to {fname}, but that shard does not hold it")

    def _resolve(self, source):
        self._index_map = None
        if isinstance(source, (list, tuple)):
            return [os.fspath(p) for p in source]
        source = os.fspath(source)
        if os.path.isfile(source):
            return [source]
        single = os.path.join(source, "model.safetensors")
        if os.path.exists(single):
            return [single]
        index = os.path.join(source, "model.safetensors.index.json")
        if os.path.exists(index):
            with open(index) as f:
                wm = json.load(f).get("weight_map")
            if not isinstance(wm, dict):
                raise W4A16Error("missing weight_map in index")   # loader.rs:431-433 wording
            self._index_map = {str(k): str(v) for k, v in wm.items()}
            names = sorted(set(self._index_map.values()), key=_natural_key)
            return [os.path.join(source, n) for n in names]
        if os.path.isdir(source):
            return [os.path.join(source, n) for n in sorted(os.listdir(source), key=_natural_key)
                    if n.endswith(".safetensors")]
        return []

    # Mapping protocol
    def __getitem__(self, name: str) -> torch.Tensor:
        f = self._where.get(name)
        if f is None:
            raise KeyError(name)
        return f.tensor(name)

    def __iter__(self) -> Iterator[str]:
        return iter(self._where)

    def __len__(self) -> int:
        return len(self._where)

    def __contains__(self, name) -> bool:
        return name in self._where

    def shape(self, name: str):
        return tuple(self._where[name].entries[name]["shape"])

    def dtype_tag(self, name: str) -> str:
        return self._where[name].entries[name]["dtype"]

    def fingerprint(self) -> str:
        """Identity of the checkpoint for the W4T cache: the hash of every file's header (names, dtypes, shapes,
        offsets) and its size -- cheap, and it changes whenever a tensor is added, resized or moved."""
        return "|".join(f.header_hash for f in self._files)

    def quantized_prefixes(self):
        """Module prefixes that hold a packed linear ("<prefix>.qweight" present)."""
        return sorted(n[: -len(".qweight")] for n in self._where if n.endswith(".qweight"))

    def dense_prefixes(self):
        """Module prefixes with a plain "<prefix>.weight" (embeddings, norms, lm_head, modules_to_not_convert)."""
        return sorted(n[: -len(".weight")] for n in self._where if n.endswith(".weight"))


def _natural_key(name: str):
    out, num = [], ""
    for ch in name:
        if ch.isdigit():
            num += ch
        else:
            if num:
                out.append((0, int(num), ""))
                num = ""
            out.append((1, 0, ch))
    if num:
        out.append((0, int(num), ""))
    return out


def save_safetensors(tensors: Mapping, path, metadata: Optional[Mapping] = None) -> None:
    """Write `tensors` (name -> CPU torch.Tensor) in safetensors layout: sorted names, 8-byte aligned header."""
    header, blobs, off = {}, [], 0
    if metadata:
        header["__metadata__"] = {str(k): str(v) for k, v in metadata.items()}
    for name in sorted(tensors):
        t = tensors[name].detach().cpu().contiguous()
        if t.dtype not in _TAG_OF:
            raise W4A16Error(f"save_safetensors: dtype {t.dtype} of '{name}' not supported")
        raw = t.view(torch.uint8).numpy().tobytes() if t.dtype is not torch.bool else t.to(torch.uint8).numpy().tobytes()
        header[name] = {"dtype": _TAG_OF[t.dtype], "shape": list(t.shape), "data_offsets": [off, off + len(raw)]}
        blobs.append(raw)
        off += len(raw)
    hj = json.dumps(header, separators=(",", ":")).encode("utf-8")
    hj += b" " * (-len(hj) % 8)
    tmp = os.fspath(path) + ".tmp"
    with open(tmp, "wb") as f:
        f.write(struct.pack("<Q", len(hj)))
        f.write(hj)
        for b in blobs:
            f.write(b)
    os.replace(tmp, path)


def save_sharded(tensors: Mapping, directory, shards: int) -> None:
    """Split `tensors` over `shards` files + model.safetensors.index.json (round-robin by sorted name)."""
    os.makedirs(directory, exist_ok=True)
    names = sorted(tensors)
    weight_map = {}
    for i in range(shards):
        fname = f"model-{i + 1:05d}-of-{shards:05d}.safetensors"
        part = {n: tensors[n] for n in names[i::shards]}
        save_safetensors(part, os.path.join(directory, fname))
        weight_map.update({n: fname for n in part})
    with open(os.path.join(directory, "model.safetensors.index.json"), "w") as f:
        json.dump({"metadata": {}, "weight_map": weight_map}, f)


# -- checkpoint -> resident layout, one decoder layer at a time --------------------------------------------------------
_SIBLINGS = {   # fused handle -> the reference's separate linears (qwen3_quantized.rs:84-96, 501-520)
    "qkv": ("self_attn.q_proj", "self_attn.k_proj", "self_attn.v_proj"),
    "gate_up": ("mlp.gate_proj", "mlp.up_proj"),
}


def load_decoder_linears(model_dir, device, layers=None, cache_dir=None, fuse: bool = True, swiglu: bool = True):
    """Every W4A16 linear of a Qwen3 / Llama style checkpoint directory, repacked (or read from the W4T cache):
    returns ({layer index: {name: QuantizedLinear}}, {prefix: dense tensor}, the weight loader).

    config.json's quantization_config picks the method (detection.rs:24-78); per layer the reference loads seven
    linears (qwen3_quantized.rs:84-96, 501-520); with `fuse` q/k/v become one handle "qkv" and gate/up one handle
    "gate_up" (SwiGLU epilogue when `swiglu`), else the seven names of the reference.  A module listed in
    modules_to_not_convert (awq.rs:243-251) is NOT on the W4A16 path: its plain ".weight" comes back in the second
    dict (the caller's dense linear takes it), and a fused group with a skipped member is loaded unfused."""
    from .quantization import EPILOGUE_NONE, EPILOGUE_SWIGLU, create_weight_loader_with_params, detect_from_directory
    model_dir = os.fspath(model_dir)
    detected = detect_from_directory(model_dir)
    with open(os.path.join(model_dir, "config.json")) as f:
        cfg = json.load(f)
    ckpt = SafetensorsCheckpoint(model_dir)
    loader = create_weight_loader_with_params(ckpt, detected, device, cache_dir)
    H, I = int(cfg["hidden_size"]), int(cfg["intermediate_size"])
    heads, kv_heads = int(cfg["num_attention_heads"]), int(cfg.get("num_key_value_heads", cfg["num_attention_heads"]))
    hd = int(cfg.get("head_dim", H // heads))
    bias = bool(cfg.get("attention_bias", False))
    shapes = {"self_attn.q_proj": (H, heads * hd, bias), "self_attn.k_proj": (H, kv_heads * hd, bias),
              "self_attn.v_proj": (H, kv_heads * hd, bias), "self_attn.o_proj": (heads * hd, H, False),
              "mlp.gate_proj": (H, I, False), "mlp.up_proj": (H, I, False), "mlp.down_proj": (I, H, False)}
    n_layers = int(cfg["num_hidden_layers"])
    out, dense = {}, {}
    skipped = loader.config.is_layer_skipped

    def one(prefix, mod):
        k, n, b = shapes[mod]
        if skipped(prefix):
            dense[prefix] = ckpt[f"{prefix}.weight"]
            return None
        return loader.load_linear(prefix, k, n, b)

    for li in (range(n_layers) if layers is None else layers):
        base = f"model.layers.{li}"
        lay = {}
        done = set()
        if fuse:
            for fused, mods in _SIBLINGS.items():
                prefixes = [f"{base}.{m}" for m in mods]
                if any(skipped(p) for p in prefixes):
                    continue
                ep = EPILOGUE_SWIGLU if (fused == "gate_up" and swiglu) else EPILOGUE_NONE
                lay[fused] = loader.load_fused(prefixes, shapes[mods[0]][0], [shapes[m][1] for m in mods],
                                               shapes[mods[0]][2], ep)
                done.update(mods)
        for mod in shapes:
            if mod in done:
                continue
            lin = one(f"{base}.{mod}", mod)
            if lin is not None:
                lay[mod.split(".")[-1]] = lin
        out[li] = lay
    return out, dense, loader
