"""Model materialisation from HF-layout safetensors and training checkpoints.

Reference: ``scaletorch/utils/checkpoint.py:23-560``.

Cold start -- same flow as the reference (meta-device construction → per-rank list of needed HF tensor
names → read from ``model.safetensors`` / sharded index → rename → TP-slice → assign), with three
differences: (1) loaded weights are KEPT (the reference calls ``reset_parameters()`` afterwards and
overwrites them); (2) the framework stores fused ``qkv_proj`` / ``gate_up_proj`` and stacked
expert tensors, so the loader assembles them from the HF ``q/k/v_proj``, ``gate/up_proj`` and
``experts.N.*`` tensors -- and ``export_hf_state_dict`` splits them back, which the reference cannot do;
(3) tied embeddings stay tied under TP/PP.

Training checkpoints -- one file per (tp, pp, ep) coordinate (the reference omits ep and lets EP ranks
clobber each other, §2.7 #9), written by dp_rank = cp_rank = 0: model, optimiser (flat fp32 state),
trained steps/tokens, RNG state and data-loader position.
"""
from __future__ import annotations

import contextlib
import json
import os
import re
from typing import Any

import torch
import torch.nn as nn

from ..parallel.process_group import process_group_manager as pgm
from .logger_utils import get_logger

logger = get_logger(__name__)


# ------------------------------------------------------------------ meta-device construction
@contextlib.contextmanager
def init_model_with_dematerialized_weights(include_buffers: bool = False):
    """Parameters created inside the context live on the ``meta`` device (no memory, no init cost)."""
    old_register = nn.Module.register_parameter
    old_buffer = nn.Module.register_buffer

    def register_parameter(module, name, param):
        old_register(module, name, param)
        if param is not None and param.device.type != "meta":
            kwargs = dict(module._parameters[name].__dict__)
            kwargs["requires_grad"] = param.requires_grad
            cls = type(module._parameters[name])
            module._parameters[name] = cls(module._parameters[name].to(torch.device("meta")), **kwargs)

    def register_buffer(module, name, buffer, persistent=True):
        old_buffer(module, name, buffer, persistent=persistent)
        if buffer is not None and include_buffers:
            module._buffers[name] = module._buffers[name].to(torch.device("meta"))

    try:
        nn.Module.register_parameter = register_parameter
        nn.Module.register_buffer = register_buffer
        with torch.device("meta"):
            yield
    finally:
        nn.Module.register_parameter = old_register
        nn.Module.register_buffer = old_buffer


def _find_safetensors(save_dir: str) -> tuple[str | None, dict | None]:
    single = os.path.join(save_dir, "model.safetensors")
    index = os.path.join(save_dir, "model.safetensors.index.json")
    if os.path.exists(single):
        return single, None
    if os.path.exists(index):
        with open(index) as f:
            return None, json.load(f)["weight_map"]
    return None, None


def _rebuild_rope(model: nn.Module) -> None:
    """RoPE tables are plain buffers computed in ``__init__``; ``to_empty`` re-allocates every buffer
    uninitialised (and the meta context may have produced meta tensors), so they are always recomputed."""
    for m in model.modules():
        if isinstance(m.__dict__.get("_buffers", {}).get("cos"), torch.Tensor) and hasattr(m, "head_dim"):
            from ..ops.norm import get_cos_sin

            cfg = m.config
            theta = getattr(cfg, "rope_theta", None)
            if theta is None:
                rp = getattr(cfg, "rope_parameters", None) or {}
                theta = rp.get("rope_theta", m.default_rope_theta) if isinstance(rp, dict) else m.default_rope_theta
            cos, sin = get_cos_sin(cfg.max_position_embeddings, m.head_dim, float(theta), device="cpu")
            m.cos, m.sin = cos, sin


def _to_empty(model: nn.Module, device: str | torch.device) -> None:
    """``Module.to_empty`` with plain ``torch.empty`` per tensor.  ``to_empty`` calls ``empty_like`` on the META tensor, which
    runs the Python reference implementation and imports the symbolic-shapes machinery: about three seconds on every rank
    of every launch for uninitialised storage of the same shape and dtype."""
    dev = torch.device(device)
    model._apply(lambda t: torch.empty(t.shape, dtype=t.dtype, device=dev))


def init_model_with_materialized_weights(model: nn.Module, model_config: Any, save_dir: str | None = None,
                                         load_weights: bool = False, device: str | torch.device = "cpu") -> nn.Module:
    """Materialises a meta-constructed model on ``device``.

    No safetensors (or ``load_weights=False``) → ``to_empty`` + ``reset_parameters()`` (the path every
    random-init benchmark takes, reference ``checkpoint.py:90-97``).  Otherwise the HF tensors this rank
    needs are read, renamed, TP/EP-sliced, fused and assigned -- and kept.
    """
    single, weight_map = _find_safetensors(save_dir) if save_dir else (None, None)
    tied = _tied_pairs(model)
    _to_empty(model, device)
    _retie(model, tied)
    _rebuild_rope(model)
    if not load_weights or (single is None and weight_map is None):
        model.reset_parameters()
        return model
    manager = InitializationManager(model, model_config)
    state = manager.load_hf_state(save_dir, single, weight_map)
    missing = [k for k in dict(model.named_parameters()) if k not in state]
    strict = not missing
    model.load_state_dict(state, strict=False)
    if not strict:
        logger.warning("HF checkpoint did not provide %d tensors (e.g. %s); they keep their random init",
                       len(missing), missing[:3])
        fresh = dict(model.named_parameters())
        saved = {k: fresh[k].detach().clone() for k in state if k in fresh}
        model.reset_parameters()
        with torch.no_grad():
            for k, v in saved.items():
                fresh[k].copy_(v)
    _retie(model, tied)
    return model


def _tied_pairs(model: nn.Module) -> list[tuple[str, str]]:
    seen: dict[int, str] = {}
    pairs = []
    for name, p in model.named_parameters(remove_duplicate=False):
        if id(p) in seen:
            pairs.append((seen[id(p)], name))
        else:
            seen[id(p)] = name
    return pairs


def _get_module_attr(model: nn.Module, dotted: str):
    mod_name, _, attr = dotted.rpartition(".")
    return (model.get_submodule(mod_name) if mod_name else model), attr


def _retie(model: nn.Module, pairs: list[tuple[str, str]]) -> None:
    for src, dst in pairs:
        sm, sa = _get_module_attr(model, src)
        dm, da = _get_module_attr(model, dst)
        dm._parameters[da] = sm._parameters[sa]


# ------------------------------------------------------------------ HF <-> local naming
class InitializationManager:
    """Maps between the HF checkpoint layout and this framework's (fused, sharded) parameters."""

    def __init__(self, model: nn.Module, model_config: Any):
        self.model = model
        self.cfg = model_config
        self.tp = pgm.tp_world_size if pgm else 1
        self.tp_rank = pgm.tp_rank if pgm else 0
        self.ep = pgm.ep_world_size if pgm else 1
        self.ep_rank = pgm.ep_rank if pgm else 0
        self.model_type = getattr(model_config, "model_type", "llama")
        self.head_dim = getattr(model_config, "head_dim", None) or model_config.hidden_size // model_config.num_attention_heads
        # global layer index of each local decoder layer (pipeline stages hold a slice)
        self.layer_ids = list(getattr(model, "layer_ids", range(model_config.num_hidden_layers)))

    # -- which HF tensors does this rank need? (reference ``checkpoint.py:265-337``)
    def hf_names_for_local(self, local_name: str) -> list[str]:
        m = re.match(r"decoder_layers\.(\d+)\.(.*)", local_name)
        if m:
            li = self.layer_ids[int(m.group(1))] if int(m.group(1)) < len(self.layer_ids) else int(m.group(1))
            rest = m.group(2)
            p = f"model.layers.{li}."
            table = {
                "input_layernorm.weight": [p + "input_layernorm.weight"],
                "post_attention_layernorm.weight": [p + "post_attention_layernorm.weight"],
                "attention.qkv_proj.weight": [p + f"self_attn.{x}_proj.weight" for x in "qkv"],
                "attention.qkv_proj.bias": [p + f"self_attn.{x}_proj.bias" for x in "qkv"],
                "attention.out_proj.weight": [p + "self_attn.o_proj.weight"],
                "attention.q_norm.weight": [p + "self_attn.q_norm.weight"],
                "attention.k_norm.weight": [p + "self_attn.k_norm.weight"],
                "mlp.gate_up_proj.weight": [p + "mlp.gate_proj.weight", p + "mlp.up_proj.weight"],
                "mlp.down_proj.weight": [p + "mlp.down_proj.weight"],
                "moe.router.gate.weight": [p + "mlp.gate.weight"],
            }
            if rest in table:
                return table[rest]
            e_local = getattr(self.cfg, "num_experts", 0) // self.ep if getattr(self.cfg, "num_experts", 0) else 0
            experts = range(self.ep_rank * e_local, (self.ep_rank + 1) * e_local)
            if rest == "moe.experts.gate_up_proj":
                return [p + f"mlp.experts.{e}.{x}_proj.weight" for e in experts for x in ("gate", "up")]
            if rest == "moe.experts.down_proj":
                return [p + f"mlp.experts.{e}.down_proj.weight" for e in experts]
            raise KeyError(f"no HF mapping for parameter '{local_name}'")
        top = {"embedding.weight": ["model.embed_tokens.weight"], "final_norm.weight": ["model.norm.weight"],
               "final_proj.weight": ["lm_head.weight"]}
        if local_name in top:
            return top[local_name]
        raise KeyError(f"no HF mapping for parameter '{local_name}'")

    # -- TP slicing helpers (reference ``checkpoint.py:339-423``)
    def _rows(self, t: torch.Tensor) -> torch.Tensor:
        n = t.shape[0] // self.tp
        return t[self.tp_rank * n:(self.tp_rank + 1) * n]

    def _cols(self, t: torch.Tensor) -> torch.Tensor:
        n = t.shape[1] // self.tp
        return t[:, self.tp_rank * n:(self.tp_rank + 1) * n]

    def assemble(self, local_name: str, hf: dict[str, torch.Tensor]) -> torch.Tensor:
        names = self.hf_names_for_local(local_name)
        ts = [hf[n] for n in names]
        if local_name.endswith(("qkv_proj.weight", "qkv_proj.bias", "gate_up_proj.weight")) and "experts" not in local_name:
            return torch.cat([self._rows(t) for t in ts], dim=0)          # per-rank [q_r; k_r; v_r] / [gate_r; up_r]
        if local_name.endswith(("out_proj.weight", "mlp.down_proj.weight")):
            return self._cols(ts[0])
        if local_name in ("embedding.weight", "final_proj.weight"):
            return self._rows(ts[0])
        if local_name.endswith("moe.experts.gate_up_proj"):
            return torch.stack([torch.cat([ts[2 * i], ts[2 * i + 1]], dim=0) for i in range(len(ts) // 2)])
        if local_name.endswith("moe.experts.down_proj"):
            return torch.stack(ts)
        return ts[0]                                                      # norms, router: replicated

    def load_hf_state(self, save_dir: str, single: str | None, weight_map: dict | None) -> dict[str, torch.Tensor]:
        from safetensors import safe_open

        local_params = dict(self.model.named_parameters())
        tied_lm_head = getattr(self.cfg, "tie_word_embeddings", False)
        needed: dict[str, list[str]] = {}
        for name in local_params:
            try:
                needed[name] = self.hf_names_for_local(name)
            except KeyError:
                logger.warning("parameter %s has no HF counterpart", name)
        want = {n for names in needed.values() for n in names}
        files: dict[str, list[str]] = {}
        for n in want:
            if weight_map is not None:
                f = weight_map.get(n)
                if f is None and n == "lm_head.weight" and tied_lm_head:
                    continue
                if f is None:
                    continue
                files.setdefault(os.path.join(save_dir, f), []).append(n)
            else:
                files.setdefault(single, []).append(n)
        hf: dict[str, torch.Tensor] = {}
        for path, names in files.items():
            with safe_open(path, framework="pt", device="cpu") as f:
                keys = set(f.keys())
                for n in names:
                    if n in keys:
                        hf[n] = f.get_tensor(n)
        if tied_lm_head and "lm_head.weight" not in hf and "model.embed_tokens.weight" in hf:
            hf["lm_head.weight"] = hf["model.embed_tokens.weight"]
        state = {}
        for name, names in needed.items():
            if all(n in hf for n in names):
                t = self.assemble(name, hf)
                if tuple(t.shape) != tuple(local_params[name].shape):
                    raise ValueError(f"shape mismatch for {name}: checkpoint {tuple(t.shape)} vs model {tuple(local_params[name].shape)}")
                state[name] = t.to(local_params[name].dtype)
        return state


def hf_state_from_unsharded(state: dict[str, torch.Tensor], model_config: Any) -> dict[str, torch.Tensor]:
    """Splits the fused tensors of an unsharded (tp = pp = ep = 1) state dict of this framework into HF tensor names --
    the inverse of :meth:`InitializationManager.assemble`.  Works on plain state dicts, so merged training checkpoints
    (``utils/sharding.py::merge_shards``, ``tools/convert_checkpoint.py``) can be exported without building a model."""
    from types import SimpleNamespace

    mgr = InitializationManager.__new__(InitializationManager)
    mgr.model, mgr.cfg = SimpleNamespace(), model_config
    mgr.tp = mgr.ep = 1
    mgr.tp_rank = mgr.ep_rank = 0
    mgr.model_type = getattr(model_config, "model_type", "llama")
    mgr.head_dim = getattr(model_config, "head_dim", None) or model_config.hidden_size // model_config.num_attention_heads
    mgr.layer_ids = list(range(model_config.num_hidden_layers))
    out: dict[str, torch.Tensor] = {}
    hd = mgr.head_dim
    nq, nkv = model_config.num_attention_heads, model_config.num_key_value_heads
    tied = getattr(model_config, "tie_word_embeddings", False)
    for name, p in state.items():
        if tied and name == "final_proj.weight":
            continue
        try:
            names = mgr.hf_names_for_local(name)
        except KeyError:
            continue                                                          # buffers (RoPE tables) have no HF counterpart
        t = p.detach().cpu()
        if name.endswith(("qkv_proj.weight", "qkv_proj.bias")):
            parts = torch.split(t, [nq * hd, nkv * hd, nkv * hd], dim=0)
        elif name.endswith("mlp.gate_up_proj.weight"):
            parts = torch.chunk(t, 2, dim=0)
        elif name.endswith("moe.experts.gate_up_proj"):
            parts = [x for e in range(t.shape[0]) for x in torch.chunk(t[e], 2, dim=0)]
        elif name.endswith("moe.experts.down_proj"):
            parts = [t[e] for e in range(t.shape[0])]
        else:
            parts = [t]
        for n, part in zip(names, parts):
            out[n] = part.contiguous()
    if tied:
        out.pop("lm_head.weight", None)
    return out


def export_hf_state_dict(model: nn.Module, model_config: Any) -> dict[str, torch.Tensor]:
    """Splits the fused parameters of an *unsharded* (tp = ep = 1) model back into HF tensor names."""
    if pgm and (pgm.tp_world_size > 1 or pgm.ep_world_size > 1):
        raise RuntimeError("export_hf_state_dict needs an unsharded model (tp = ep = 1); merge sharded checkpoints "
                           "with tools/convert_checkpoint.py")
    return hf_state_from_unsharded(dict(model.named_parameters()), model_config)


# ------------------------------------------------------------------ training checkpoints
def find_complete_checkpoints(work_dir: str) -> list[tuple[int, str]]:
    """``[(step, path)]`` of the checkpoint directories ``work_dir/<step>`` that hold a full set of shard files, oldest
    first.  A directory with a ``*.tmp`` file (a writer was interrupted) or with fewer ``weights_*.pth`` files than the
    fullest one (some ranks had not written yet when the job died) does not count."""
    found = []
    if not os.path.isdir(work_dir):
        return []
    for name in os.listdir(work_dir):
        path = os.path.join(work_dir, name)
        if name.isdigit() and os.path.isdir(path):
            files = os.listdir(path)
            shards = [f for f in files if f.startswith("weights_") and f.endswith(".pth")]
            if shards and not any(f.endswith(".tmp") for f in files):
                found.append((int(name), path, len(shards)))
    if not found:
        return []
    found.sort()
    full = max(n for _, _, n in found)
    return [(s, p) for s, p, n in found if n == full]


def resolve_resume_path(resume_path: str, work_dir: str) -> str:
    """``--resume_path latest`` (or ``auto``) -> the newest complete checkpoint under ``work_dir`` (``""`` = start from
    scratch when there is none); any other value is returned unchanged."""
    if resume_path not in ("latest", "auto"):
        return resume_path
    complete = find_complete_checkpoints(work_dir)
    return complete[-1][1] if complete else ""


def prune_checkpoints(work_dir: str, keep: int, protect: str | None = None) -> list[str]:
    """Delete all but the ``keep`` newest COMPLETE checkpoints under ``work_dir`` (``keep <= 0``: keep everything).
    ``protect`` (the directory being written right now) is never touched, and incomplete directories older than the newest
    complete one are removed as debris.  Returns the deleted paths.  Call it from ONE process (global rank 0)."""
    import shutil

    if keep <= 0:
        return []
    complete = find_complete_checkpoints(work_dir)
    doomed = [p for _, p in complete[:-keep]] if len(complete) > keep else []
    if complete and os.path.isdir(work_dir):
        newest = complete[-1][0]
        keep_paths = {p for _, p in complete}
        for name in os.listdir(work_dir):
            path = os.path.join(work_dir, name)
            if name.isdigit() and os.path.isdir(path) and path not in keep_paths and int(name) < newest:
                doomed.append(path)
    protect = os.path.abspath(protect) if protect else None
    deleted = []
    for path in doomed:
        if protect and os.path.abspath(path) == protect:
            continue
        shutil.rmtree(path, ignore_errors=True)
        deleted.append(path)
    return deleted


_SHARD_NAME = re.compile(r"weights_tp_rank_world_size=(\d+)_(\d+)_pp_rank_world_size=(\d+)_(\d+)(?:_ep_rank_world_size=(\d+)_(\d+))?\.pth$")


def find_shards(ckpt_dir: str) -> tuple[dict[tuple[int, int, int], str], tuple[int, int, int]]:
    """``{(tp_rank, pp_rank, ep_rank): path}`` and the layout ``(tp, pp, ep)`` found in ``ckpt_dir``."""
    found: dict[tuple[int, int, int], str] = {}
    layout = None
    for f in sorted(os.listdir(ckpt_dir)):
        m = _SHARD_NAME.match(f)
        if not m:
            continue
        tr, tp, pr, pp = (int(m.group(i)) for i in range(1, 5))
        er, ep = (int(m.group(5)), int(m.group(6))) if m.group(5) is not None else (0, 1)
        if layout is None:
            layout = (tp, pp, ep)
        elif layout != (tp, pp, ep):
            raise ValueError(f"{ckpt_dir} mixes layouts {layout} and {(tp, pp, ep)}")
        found[(tr, pr, er)] = os.path.join(ckpt_dir, f)
    if layout is None:
        raise FileNotFoundError(f"no weights_tp_rank_world_size=*.pth files in {ckpt_dir}")
    return found, layout


def merge_checkpoint(ckpt_dir: str, model_config) -> tuple[dict[str, torch.Tensor], dict]:
    """Returns ``(unsharded model state dict, meta)``; ``meta`` carries the step / token counters and the source layout."""
    from .sharding import merge_shards

    paths, (tp, pp, ep) = find_shards(ckpt_dir)
    want = tp * pp * ep
    if len(paths) != want:
        raise FileNotFoundError(f"{ckpt_dir} holds {len(paths)} of the {want} shard files of its tp{tp} pp{pp} ep{ep} layout")
    states, meta = {}, {}
    moments: dict[str, dict] | None = {"exp_avg": {}, "exp_avg_sq": {}}
    optimizer_step = 0
    for coord, path in paths.items():
        blob = torch.load(path, map_location="cpu", weights_only=False)
        states[coord] = blob["model"]
        opt, slots = blob.get("optimizer"), blob.get("optimizer_slots")
        if moments is not None and slots and isinstance(opt, dict) and torch.is_tensor(opt.get("exp_avg")):
            for key in ("exp_avg", "exp_avg_sq"):
                flat = opt[key]
                moments[key][coord] = {n: flat[o:o + k].view(shape) for n, o, k, shape in slots}
            optimizer_step = int(opt.get("step", 0))
        else:
            moments = None                              # a shard without a slot table: the moments cannot be re-cut
        meta = {"trained_steps": int(blob.get("trained_steps", 0)), "trained_tokens": int(blob.get("trained_tokens", 0))}
        dl = (blob.get("extra") or {}).get("data_loader")
        if dl:
            meta["data_loader"] = dl                    # position of the sample stream (layout-independent form inside)
    full = merge_shards(states, model_config, tp, pp, ep)
    meta["source_layout"] = {"tp": tp, "pp": pp, "ep": ep}
    if moments is not None:
        # AdamW moments have the shapes of their parameters: the same merge rules apply, name by name
        meta["moments"] = {key: merge_shards(moments[key], model_config, tp, pp, ep) for key in moments}
        meta["optimizer_step"] = optimizer_step
    return full, meta


class CheckpointManager:
    """Per-(tp, pp, ep) ``torch.save`` checkpoints under ``work_dir/<step>/`` (reference ``checkpoint.py:467-560``)."""

    def __init__(self, async_save: bool = False) -> None:
        """``async_save``: the step returns as soon as the state has been copied to host memory; serialisation and the
        atomic publish run on a background thread (one save in flight: the next save -- or :meth:`wait` -- joins it).
        The reference writes synchronously from the training loop (``checkpoint.py:467-520``)."""
        self.async_save = bool(async_save)
        self._writer = None              # background thread of the save in flight
        self._writer_error: BaseException | None = None
        self.tp_rank = pgm.tp_rank if pgm else 0
        self.tp_size = pgm.tp_world_size if pgm else 1
        self.pp_rank = pgm.pp_rank if pgm else 0
        self.pp_size = pgm.pp_world_size if pgm else 1
        self.ep_rank = pgm.ep_rank if pgm else 0
        self.ep_size = pgm.ep_world_size if pgm else 1
        self.cp_dp_first = (pgm.dp_rank == 0 and pgm.cp_rank == 0) if pgm else True

    def _filename(self) -> str:
        name = (f"weights_tp_rank_world_size={self.tp_rank}_{self.tp_size}"
                f"_pp_rank_world_size={self.pp_rank}_{self.pp_size}")
        if self.ep_size > 1:
            name += f"_ep_rank_world_size={self.ep_rank}_{self.ep_size}"
        return name + ".pth"

    @staticmethod
    def _unwrap(model: nn.Module) -> nn.Module:
        while hasattr(model, "module") and isinstance(model.module, nn.Module):
            model = model.module
        return model

    def save_checkpoint(self, model: nn.Module, optimizer, trained_steps: int, trained_tokens: int, out_dir: str,
                        extra: dict | None = None) -> str | None:
        # a sharded (ZeRO-1) optimiser assembles its state collectively: every rank must take part before the
        # non-writing ranks leave
        sharded = optimizer is not None and (hasattr(optimizer, "shards") or hasattr(optimizer, "sync_params"))
        opt_state = optimizer.state_dict() if sharded else None      # (also completes a deferred parameter gather)
        if not self.cp_dp_first:
            return None
        os.makedirs(out_dir, exist_ok=True)
        path = os.path.join(out_dir, self._filename())
        raw = self._unwrap(model)
        payload = {
            # a host SNAPSHOT: with async_save the writer thread must not see later updates (CPU tensors are cloned)
            "model": _to_cpu(dict(raw.state_dict()), copy=self.async_save),
            "optimizer": (_to_cpu(opt_state if opt_state is not None else optimizer.state_dict(), copy=self.async_save)
                          if optimizer is not None else None),
            "trained_steps": int(trained_steps),
            "trained_tokens": int(trained_tokens),
            "rng": {"torch": torch.get_rng_state(),
                    "cuda": torch.cuda.get_rng_state() if torch.cuda.is_available() else None},
            "layout": {"tp": self.tp_size, "pp": self.pp_size, "ep": self.ep_size},
        }
        store = getattr(optimizer, "store", None)
        if store is not None and hasattr(store, "slots"):
            # where each parameter's moments sit in the flat exp_avg / exp_avg_sq: lets another layout re-cut them by name
            payload["optimizer_slots"] = [(sl.name, int(sl.offset), int(sl.numel), tuple(sl.param.shape)) for sl in store.slots]
        if extra:
            payload["extra"] = extra
        self.wait()                                   # at most one save in flight (and surface an earlier failure)
        if self.async_save:
            import threading

            self._writer = threading.Thread(target=self._write, args=(payload, path), name="stb-checkpoint-writer", daemon=False)
            self._writer.start()
        else:
            self._write(payload, path)
            self._raise_writer_error()
        return path

    def _write(self, payload: dict, path: str) -> None:
        try:
            tmp = path + ".tmp"
            torch.save(payload, tmp)
            os.replace(tmp, path)  # atomic publish: a crash never leaves a truncated checkpoint behind
        except BaseException as e:  # reported by the next save / wait(), never lost
            self._writer_error = e

    def _raise_writer_error(self) -> None:
        if self._writer_error is not None:
            e, self._writer_error = self._writer_error, None
            raise RuntimeError(f"checkpoint write failed: {type(e).__name__}: {e}") from e

    def wait(self) -> None:
        """Blocks until the save in flight (if any) has been published; raises if it failed."""
        if self._writer is not None:
            self._writer.join()
            self._writer = None
        self._raise_writer_error()

    def layout_matches(self, load_dir: str) -> bool:
        """Whether ``load_dir`` was written under this run's (tp, pp, ep) layout."""
        _, layout = find_shards(load_dir)
        return layout == (self.tp_size, self.pp_size, self.ep_size)

    def load_resharded(self, model: nn.Module, load_dir: str, model_config, optimizer=None) -> tuple[int, int]:
        """Resume the WEIGHTS of a checkpoint written under another (tp, pp, ep) layout: every rank merges the shard files
        into the unsharded state (``merge_checkpoint``) and cuts its own slice (``load_from_unsharded``), pipeline stage
        included.  Step and token counters and the data position (kept in samples, hence layout-independent) carry over.
        The AdamW moments have the shapes of their parameters, so when the checkpoint carries the slot table of its flat
        optimiser state (``optimizer_slots``, written since this feature exists) they are merged and re-cut by the same
        rules and loaded into ``optimizer`` (``self.moments_resharded``); otherwise they restart from zero, which the caller
        must announce.  The reference cannot resume across layouts at all (``checkpoint.py:467-560``)."""
        from .sharding import load_from_unsharded

        full, meta = merge_checkpoint(load_dir, model_config)
        raw = self._unwrap(model)
        load_from_unsharded(raw, full, model_config, tp=self.tp_size, tp_rank=self.tp_rank, ep=self.ep_size, ep_rank=self.ep_rank,
                            layer_ids=getattr(raw, "layer_ids", None))
        self.last_extra = {"data_loader": meta["data_loader"]} if "data_loader" in meta else None
        self.last_source_layout = meta["source_layout"]
        self.moments_resharded = False
        store = getattr(optimizer, "store", None)
        if optimizer is not None and "moments" in meta and store is not None and hasattr(store, "slots"):
            from .sharding import shard_tensor

            layer_ids = getattr(raw, "layer_ids", None)
            sd = {"step": meta["optimizer_step"], "param_groups": []}
            for key in ("exp_avg", "exp_avg_sq"):
                flat = torch.zeros(store.total_numel, dtype=torch.float32)
                for sl in store.slots:
                    m = re.match(r"decoder_layers\.(\d+)\.(.*)", sl.name)
                    src = f"decoder_layers.{layer_ids[int(m.group(1))]}.{m.group(2)}" if (m and layer_ids is not None) else sl.name
                    if src not in meta["moments"][key] and src == "final_proj.weight":
                        src = "embedding.weight"           # tied embeddings
                    piece = shard_tensor(src, meta["moments"][key][src], model_config, self.tp_size, self.tp_rank, self.ep_size, self.ep_rank)
                    flat[sl.offset:sl.offset + sl.numel] = piece.reshape(-1).float()
                sd[key] = flat
            optimizer.load_state_dict(sd)
            self.moments_resharded = True
        return meta["trained_steps"], meta["trained_tokens"]

    def load_checkpoint(self, model: nn.Module, optimizer, load_dir: str, restore_rng: bool = True) -> tuple[int, int]:
        path = os.path.join(load_dir, self._filename())
        if not os.path.exists(path):
            raise FileNotFoundError(f"checkpoint {path} not found (checkpoints must be resumed with the same TP/PP/EP layout)")
        ckpt = torch.load(path, map_location="cpu", weights_only=False)
        lay = ckpt.get("layout")
        if lay and (lay["tp"], lay["pp"], lay["ep"]) != (self.tp_size, self.pp_size, self.ep_size):
            raise ValueError(f"checkpoint layout {lay} differs from the current layout")
        raw = self._unwrap(model)
        with torch.no_grad():
            own = raw.state_dict()
            for k, v in ckpt["model"].items():
                if k in own:
                    own[k].copy_(v)   # in place: parameters are views into the flat store
        if optimizer is not None and ckpt.get("optimizer") is not None:
            optimizer.load_state_dict(ckpt["optimizer"])
        if restore_rng and ckpt.get("rng"):
            torch.set_rng_state(ckpt["rng"]["torch"])
            if ckpt["rng"].get("cuda") is not None and torch.cuda.is_available():
                torch.cuda.set_rng_state(ckpt["rng"]["cuda"])
        self.last_extra = ckpt.get("extra")
        return int(ckpt["trained_steps"]), int(ckpt["trained_tokens"])


def _to_cpu(obj, copy: bool = False):
    """Tensors of a nested structure on the host; ``copy`` also clones tensors that already live there."""
    if isinstance(obj, torch.Tensor):
        t = obj.detach()
        return t.clone() if (copy and t.device.type == "cpu") else t.cpu()
    if isinstance(obj, dict):
        return {k: _to_cpu(v, copy) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_cpu(v, copy) for v in obj)
    return obj
