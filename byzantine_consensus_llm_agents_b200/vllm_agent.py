"""Drop-in replacement for the reference's `byzantine_consensus_game/vllm_agent.py`.

Register this module as `vllm_agent` and the unmodified reference —
`bcg_agents.py` (class BCGAgent(VLLMAgent), bcg_agents.py:46,134,157), `main.py`
(first_agent.batch_generate_json, main.py:305,412; VLLMAgent.shutdown, main.py:1069-1070) — runs
on the B200 engine instead of vLLM.  Same module surface as the reference file
(vllm_agent.py:31, 58-551): module global VERBOSE; class VLLMAgent with the class-level shared
engine, `generate`, `generate_json`, `batch_generate_json`, `batch_generate`,
`_format_chat_prompt`, `_extract_json`, classmethod `shutdown`; same argument meaning; same
error behaviour (never raises out of the JSON calls — returns {"error": ...} dicts,
vllm_agent.py:369-379, 441-448).

What differs is underneath: `batch_generate_json` sends ALL rows to the engine in one call
even when their schemas differ (the reference degrades to one prompt at a time,
vllm_agent.py:449-455), and there is no `max_num_seqs` concurrency cap.
"""
import gc
import io
import json
import os
import sys
from typing import Any, Callable, Dict, List, Optional, Tuple

VERBOSE = os.environ.get("VERBOSE", "0") == "1"

_DEFAULT_MODEL_CONFIG = {
    "max_model_len": 4096,
    "gpu_memory_utilization": 0.85,
    "tensor_parallel_size": 1,
    "max_num_seqs": 64,
}

# ---- chat templates (vllm_agent.py:199-292): (system, user) -> prompt string ---------------
_CHATML = "<|im_start|>system\n{s}<|im_end|>\n<|im_start|>user\n{u}<|im_end|>\n<|im_start|>assistant\n"
_CHATML_NO_THINK = ("<|im_start|>system\n{s}<|im_end|>\n<|im_start|>user\n{u} /no_think<|im_end|>\n"
                    "<|im_start|>assistant\n")
_LLAMA3 = ("<|begin_of_text|><|start_header_id|>system<|end_header_id|>\n\n{s}<|eot_id|>"
           "<|start_header_id|>user<|end_header_id|>\n\n{u}<|eot_id|>"
           "<|start_header_id|>assistant<|end_header_id|>\n\n")
_INST = "<s>[INST] <<SYS>>\n{s}\n<</SYS>>\n\n{u} [/INST]"


def _pick_template(model_name: str, disable_thinking: bool) -> Tuple[str, str]:
    m = model_name.lower()
    if "qwen3" in m or "qwen-3" in m:
        if "instruct-2507" in m or "instruct_2507" in m:
            return _CHATML, "Qwen3-Instruct-2507 template (no thinking mode)"
        if disable_thinking:
            return _CHATML_NO_THINK, "Qwen3 template with /no_think (thinking disabled)"
        return _CHATML, "Qwen3 template with thinking enabled"
    if "qwen" in m:
        return _CHATML, "Qwen2.5 ChatML template"
    if "llama-3" in m or "llama3" in m:
        return _LLAMA3, "Llama3 template"
    if "llama" in m or "mistral" in m:
        return _INST, "standard chat template"
    return _CHATML, "default ChatML template"


# ---- model registry: model name -> (ModelConfig, weights factory) --------------------------
_REGISTRY: Dict[str, Tuple[Any, Callable]] = {}


def register_model(model_name: str, cfg, weights_factory: Callable) -> None:
    """Make `model_name` loadable.  `weights_factory(cfg, device) -> Weights`.  Used by tests to
    plug in transformers-initialised weights that the CPU oracle shares."""
    _REGISTRY[model_name.lower()] = (cfg, weights_factory)


class B200LLM:
    """The object stored in `VLLMAgent._shared_llm` — stands where `vllm.LLM` stood."""

    def __init__(self, model_name: str, model_config: Dict[str, Any]):
        import torch

        from . import model as M
        from .engine import Engine
        from .tokenizer import get_tokenizer

        self.data_parallel = bool(model_config.get("data_parallel", False))
        self.tp = None
        tp_size = int(model_config.get("tensor_parallel_size", 1))
        if tp_size > 1:
            # the reference has vLLM spawn `mp` workers here (vllm_agent.py:141-142); same shape:
            # this process is rank 0, ranks 1..tp-1 are spawned, one per additional GPU
            # (or, inside a torchrun job, the replica's other ranks already exist: model_config["tp_group"])
            from .tp import GroupTP, SpawnedTP
            cls = GroupTP if model_config.get("tp_group") is not None else SpawnedTP
            self.tp = cls(model_name, model_config, tp_size)
            self.cfg, self.tokenizer, self.engine = self.tp.cfg, self.tp.tokenizer, self.tp.engine
            return
        key = model_name.lower()
        tokenizer = None
        if key in _REGISTRY:
            cfg, factory = _REGISTRY[key]
        elif os.path.isdir(model_name) and os.path.exists(os.path.join(model_name, "config.json")):
            # a local HuggingFace model directory: its own architecture, checkpoint and tokenizer —
            # what the reference gets when vLLM resolves the model name (vllm_agent.py:126-144)
            from .tokenizer import LocalTokenizer
            cfg = M.config_from_dir(model_name)
            factory = lambda c, dev: M.Weights.from_safetensors_dir(c, model_name, dev)
            if os.path.exists(os.path.join(model_name, "tokenizer.json")):
                tokenizer = LocalTokenizer(model_name, cfg.vocab, cfg.family)
        else:
            cfg = M.config_for(model_name)
            seed = int(model_config.get("weight_seed", os.environ.get("BCG_B200_WEIGHT_SEED", 0)))
            factory = lambda c, dev: M.Weights.random_init(c, dev, seed=seed)
        device = model_config.get("device", "cuda:0")
        self.cfg = cfg
        self.tokenizer = tokenizer or get_tokenizer(cfg.family, cfg.vocab)
        weights = factory(cfg, torch.device(device))
        self.engine = Engine(
            cfg, weights, self.tokenizer,
            max_model_len=int(model_config.get("max_model_len", 8192)),
            gpu_memory_utilization=float(model_config.get("gpu_memory_utilization", 0.85)),
            max_rows=int(model_config.get("engine_max_rows", 64)),
            kv_pages=model_config.get("kv_pages"),
            base_seed=int(model_config.get("seed", os.environ.get("BCG_B200_SEED", 0))),
            device=device,
            use_cuda_graph=bool(model_config.get("use_cuda_graph", True)),
            pdl=bool(model_config.get("pdl", True)),
            fused_rope=bool(model_config.get("fused_rope", False)))

    def run(self, prompts: List[str], schemas: List[Optional[dict]], temperature: float,
            max_tokens: int) -> List[str]:
        from . import dp
        from .engine import Request, row_seed
        call = self.engine.calls
        if self.tp is not None:
            reqs = [Request(p, s, float(temperature), int(max_tokens)) for p, s in zip(prompts, schemas)]
            return [r.text for r in self.tp.generate(reqs)]
        if not self.data_parallel or dp.dist_info()[1] == 1:
            reqs = [Request(p, s, float(temperature), int(max_tokens)) for p, s in zip(prompts, schemas)]
            return [r.text for r in self.engine.generate(reqs)]

        # rows of this call are split over the replicas; each row keeps the Philox key of its
        # GLOBAL index, so the gathered texts equal what one GPU would have produced
        def run_local(rows):
            reqs = [Request(prompts[i], schemas[i], float(temperature), int(max_tokens),
                            seed=row_seed(self.engine.base_seed, call, i)) for i in rows]
            return [r.text for r in self.engine.generate(reqs)]
        texts = dp.run_sharded(run_local, len(prompts))
        self.engine.calls = call + 1          # replicas with no rows still advance the call counter
        return texts

    def close(self) -> None:
        if self.tp is not None:
            self.tp.close()
        else:
            self.engine.close()


def _brace_match(text: str) -> Optional[Dict[str, Any]]:
    """First '{' ... its matching '}' that parses (vllm_agent.py:344-364, 457-472)."""
    start = text.find("{")
    if start < 0:
        return None
    depth = 0
    for i in range(start, len(text)):
        ch = text[i]
        if ch == "{":
            depth += 1
        elif ch == "}":
            depth -= 1
            if depth == 0:
                try:
                    return json.loads(text[start:i + 1])
                except ValueError:
                    continue
    return None


class VLLMAgent:
    """Base agent; all agents share one engine (class-level singleton, vllm_agent.py:65-67)."""

    _shared_llm: Optional[B200LLM] = None
    _shared_model_name: Optional[str] = None
    _shared_model_config: Optional[Dict[str, Any]] = None

    def __init__(self, agent_id: str, model_name: str = "Qwen/Qwen3-8B",
                 model_config: Optional[Dict[str, Any]] = None):
        self.agent_id = agent_id
        self.model_name = model_name
        self.model_config = model_config or dict(_DEFAULT_MODEL_CONFIG)
        cls = VLLMAgent
        if (cls._shared_llm is None or cls._shared_model_name != model_name
                or cls._shared_model_config != self.model_config):
            self._load_model()
        self.llm = cls._shared_llm

    def _load_model(self):
        if VERBOSE:
            print(f"Loading model {self.model_name}... (B200 engine, random-init weights)", flush=True)
        saved = (sys.stdout, sys.stderr)
        try:
            for k, v in self.model_config.get("env_vars", {}).items():
                os.environ[k] = v
            if not VERBOSE:
                sys.stdout, sys.stderr = io.StringIO(), io.StringIO()
            if VLLMAgent._shared_llm is not None:
                VLLMAgent.shutdown()
            VLLMAgent._shared_llm = B200LLM(self.model_name, self.model_config)
            VLLMAgent._shared_model_name = self.model_name
            VLLMAgent._shared_model_config = dict(self.model_config)
        except Exception as e:
            sys.stdout, sys.stderr = saved
            print(f"\n✗ Error loading model: {e}", flush=True)
            raise
        finally:
            sys.stdout, sys.stderr = saved
        if VERBOSE:
            print("✓ Model loaded successfully", flush=True)

    # ------------------------------------------------------------------ templates
    def _format_chat_prompt(self, system_prompt: str, user_prompt: str) -> str:
        template, label = _pick_template(self.model_name,
                                         self.model_config.get("disable_qwen3_thinking", True))
        if VERBOSE:
            print(f"Using {label}")
        return template.format(s=system_prompt, u=user_prompt)

    # ------------------------------------------------------------------ free text
    def generate(self, prompt: str, temperature: float = 0.0, max_tokens: int = 256,
                 top_p: float = 1.0, system_prompt: Optional[str] = None, **kwargs) -> str:
        self._check_sampling(top_p, kwargs)
        full = self._format_chat_prompt(system_prompt, prompt) if system_prompt else prompt
        return self.llm.run([full], [None], temperature, max_tokens)[0].strip()

    def batch_generate(self, prompts: list[str], temperature: float = 0.0, max_tokens: int = 256,
                       top_p: float = 1.0, **kwargs) -> list[str]:
        self._check_sampling(top_p, kwargs)
        if not prompts:
            return []
        return [t.strip() for t in self.llm.run(list(prompts), [None] * len(prompts), temperature,
                                                max_tokens)]

    @staticmethod
    def _check_sampling(top_p: float, kwargs: Dict[str, Any]) -> None:
        # The reference never sets top_p / extra sampling options; refusing them
        # is better than silently sampling from a different distribution.
        if top_p != 1.0 or kwargs:
            raise NotImplementedError("only temperature / max_tokens sampling is implemented "
                                      f"(got top_p={top_p}, extra={sorted(kwargs)})")

    # ------------------------------------------------------------------ guided JSON
    def generate_json(self, prompt: str, schema: Dict[str, Any], temperature: float = 0.0,
                      max_tokens: int = 512, system_prompt: Optional[str] = None) -> Dict[str, Any]:
        try:
            full = self._format_chat_prompt(system_prompt, prompt) if system_prompt else prompt
            text = self.llm.run([full], [schema], temperature, max_tokens)[0]
            return self._parse_single(text)
        except Exception as e:  # the reference never lets this call raise (vllm_agent.py:376-379)
            if VERBOSE:
                print(f"⚠️ JSON generation failed: {e}")
            return {"error": str(e), "message": "JSON generation failed"}

    @staticmethod
    def _parse_single(text: str) -> Dict[str, Any]:
        text = text.strip()
        obj = None
        try:
            obj = json.loads(text)
        except ValueError:
            obj = _brace_match(text)
        if obj is not None:
            return obj
        return {"error": "json_parse_failed",
                "message": "Could not extract valid JSON from model output", "raw": text[:200]}

    def batch_generate_json(self, prompts: List[Tuple[str, str, Dict[str, Any]]],
                            temperature: float = 0.8, max_tokens: int = 512) -> List[Dict[str, Any]]:
        if not prompts:
            return []
        formatted = [self._format_chat_prompt(s, u) for s, u, _ in prompts]
        schemas = [sc for _, _, sc in prompts]
        uniform = all(sc == schemas[0] for sc in schemas)
        try:
            texts = self.llm.run(formatted, schemas, temperature, max_tokens)
        except Exception as e:
            if VERBOSE:
                print(f"⚠️ Batched JSON generation failed: {e}")
            if uniform:
                return [{"error": str(e)} for _ in prompts]
            return [{"error": str(e), "message": "JSON generation failed"} for _ in prompts]
        results = []
        for text in texts:
            if not uniform:
                # the reference reaches generate_json() per prompt here; keep its error shape
                results.append(self._parse_single(text))
                continue
            text = text.strip()
            try:
                results.append(json.loads(text))
            except ValueError:
                obj = self._extract_json(text)
                results.append(obj if obj else {"error": "json_parse_failed", "raw": text[:100]})
        return results

    def _extract_json(self, text: str) -> Optional[Dict[str, Any]]:
        return _brace_match(text)

    # ------------------------------------------------------------------ teardown
    @classmethod
    def shutdown(cls):
        """Idempotent; safe when nothing was loaded (vllm_agent.py:506-551)."""
        llm, cls._shared_llm = cls._shared_llm, None
        cls._shared_model_name = None
        cls._shared_model_config = None
        if llm is not None:
            try:
                llm.close()
            except Exception:
                pass
            del llm
        gc.collect()
        try:
            import torch
            if torch.cuda.is_available():
                torch.cuda.synchronize()
                torch.cuda.empty_cache()
                torch.cuda.reset_peak_memory_stats()
        except Exception as e:
            if VERBOSE:
                print(f"Warning: CUDA cleanup failed: {e}", flush=True)
        # The reference also tears down torch.distributed here (vllm_agent.py:540-547) because vLLM's
        # `mp` executor had created the groups.  The counterpart of that — the control group of a
        # spawned tensor-parallel replica — is destroyed by SpawnedTP.close() above; a process group
        # the CALLER created (torchrun: data-parallel replicas, tensor-parallel replicas cut out of
        # the world) is the caller's to destroy, other ranks may still be waiting on it.
        gc.collect()
        if VERBOSE:
            print("✓ B200 engine shutdown complete", flush=True)
