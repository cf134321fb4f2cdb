"""ORACLE (test infrastructure only — never imported by the product path).

Runs the UNMODIFIED reference (`/root/reference/byzantine_consensus_game/*.py`) in this
container by supplying the third-party `vllm` package it imports (cut line B2 of SURVEY.md
§8b): `vllm.LLM`, `vllm.SamplingParams`, `vllm.sampling_params.GuidedDecodingParams`
(vllm_agent.py:13-14).  The shim forwards `LLM.generate(prompts, sampling_params)` to a
pluggable backend and records every call, which is how the golden fixtures under
tests/golden/ are produced (oracle/tools/*.py).

/root/reference exists only in the build container; nothing that runs on the GPU box
imports this module.
"""
from __future__ import annotations

import contextlib
import importlib
import io
import os
import random
import sys
import tempfile
import types
from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Callable, List, Optional

REFERENCE_DIR = Path("/root/reference/byzantine_consensus_game")
REFERENCE_MODULES = ["vllm_agent", "bcg_agents", "main", "config", "byzantine_consensus", "a2a_sim",
                     "agent_network", "communication_protocol", "protocol_factory"]


class GuidedDecodingParams:
    def __init__(self, json=None, **kw):
        self.json = json
        self.extra = kw

    def __eq__(self, other):
        return isinstance(other, GuidedDecodingParams) and self.json == other.json


class SamplingParams:
    def __init__(self, temperature=1.0, top_p=1.0, max_tokens=16, guided_decoding=None, **kw):
        self.temperature = temperature
        self.top_p = top_p
        self.max_tokens = max_tokens
        self.guided_decoding = guided_decoding
        self.extra = kw


@dataclass
class _Completion:
    text: str
    token_ids: list = field(default_factory=list)


@dataclass
class _RequestOutput:
    outputs: list


@dataclass
class CallRecord:
    prompts: List[str]
    temperature: float
    max_tokens: int
    schema: Optional[dict]
    texts: List[str]
    extra: Any = None


class ShimLLM:
    """Stands in for vllm.LLM.  `backend(prompts, sampling_params) -> list[str | (str, extra)]`."""

    backend: Callable = None
    calls: List[CallRecord] = []
    init_kwargs: List[dict] = []

    def __init__(self, **kw):
        ShimLLM.init_kwargs.append(kw)

    def generate(self, prompts, sampling_params):
        if ShimLLM.backend is None:
            raise RuntimeError("no oracle backend installed")
        res = ShimLLM.backend(list(prompts), sampling_params)
        texts, extras = [], []
        for r in res:
            if isinstance(r, tuple):
                texts.append(r[0])
                extras.append(r[1])
            else:
                texts.append(r)
                extras.append(None)
        gd = getattr(sampling_params, "guided_decoding", None)
        ShimLLM.calls.append(CallRecord(list(prompts), sampling_params.temperature,
                                        sampling_params.max_tokens, gd.json if gd else None, texts,
                                        extras))
        return [_RequestOutput([_Completion(t)]) for t in texts]


def install_vllm_shim(backend: Callable) -> None:
    ShimLLM.backend = backend
    ShimLLM.calls = []
    ShimLLM.init_kwargs = []
    vllm = types.ModuleType("vllm")
    vllm.LLM = ShimLLM
    vllm.SamplingParams = SamplingParams
    sp = types.ModuleType("vllm.sampling_params")
    sp.GuidedDecodingParams = GuidedDecodingParams
    sp.SamplingParams = SamplingParams
    vllm.sampling_params = sp
    sys.modules["vllm"] = vllm
    sys.modules["vllm.sampling_params"] = sp


def uninstall() -> None:
    for m in ["vllm", "vllm.sampling_params", *REFERENCE_MODULES]:
        sys.modules.pop(m, None)
    if str(REFERENCE_DIR) in sys.path:
        sys.path.remove(str(REFERENCE_DIR))


def import_reference(adapter_module: Optional[types.ModuleType] = None):
    """Import the reference modules fresh.  With `adapter_module`, that module is registered as
    `vllm_agent` first — i.e. the drop-in replacement of cut line B1 is exercised instead of the
    reference's own adapter."""
    if not REFERENCE_DIR.exists():
        raise RuntimeError(f"{REFERENCE_DIR} is not available on this machine")
    for m in REFERENCE_MODULES:
        sys.modules.pop(m, None)
    if adapter_module is not None:
        sys.modules["vllm_agent"] = adapter_module
    if str(REFERENCE_DIR) not in sys.path:
        sys.path.insert(0, str(REFERENCE_DIR))
    mods = {name: importlib.import_module(name) for name in REFERENCE_MODULES}
    return types.SimpleNamespace(**mods)


def run_game(ref, num_honest: int, num_byzantine: int, max_rounds: int, seed: int = 1234,
             greedy: bool = False, awareness: str = "may_exist", model_name: Optional[str] = None,
             model_config_overrides: Optional[dict] = None, quiet: bool = True,
             llm_overrides: Optional[dict] = None):
    """One BCGSimulation run of the unmodified reference in a scratch cwd.  Returns
    (statistics dict, round-by-round transcript, simulation object)."""
    cfg = ref.config
    if model_name:
        cfg.VLLM_CONFIG["model_name"] = model_name
    if model_config_overrides:
        cfg.VLLM_CONFIG.update(model_config_overrides)
    if greedy:
        cfg.LLM_CONFIG["temperature_decide"] = 0.0
        cfg.LLM_CONFIG["temperature_vote"] = 0.0
    if llm_overrides:          # e.g. shorter max_tokens for wide, CPU-expensive parity fixtures
        cfg.LLM_CONFIG.update(llm_overrides)
    cfg.METRICS_CONFIG["save_results"] = False
    random.seed(seed)
    cwd = os.getcwd()
    tmp = tempfile.mkdtemp(prefix="bcg_ref_run_")
    os.chdir(tmp)
    out = io.StringIO()
    try:
        with (contextlib.redirect_stdout(out) if quiet else contextlib.nullcontext()):
            sim = ref.main.BCGSimulation(num_honest=num_honest, num_byzantine=num_byzantine,
                                         config={"max_rounds": max_rounds,
                                                 "byzantine_awareness": awareness})
            transcript = []
            while not sim.game.game_over:
                sim.run_round()
                g = sim.game
                transcript.append({
                    "round": g.current_round,
                    "proposals": {a: s.current_value for a, s in sorted(g.agents.items())},
                    "game_over": g.game_over,
                })
            stats = sim.game.get_statistics()
    finally:
        os.chdir(cwd)
        try:
            ref.vllm_agent.VLLMAgent.shutdown()
        except Exception:
            pass
    return stats, transcript, sim
