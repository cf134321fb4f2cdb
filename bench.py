#!/usr/bin/env python
"""bench.py — per-round agent step of the Byzantine Consensus Game on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Workload (BASELINE.json configs[2], the one `metric` is quoted on): Qwen3-14B architecture,
random-init bf16 weights, 8 honest + 2 Byzantine agents.  One STEP = one game round's agent step:
every agent's *decide* call (T=0.5, <=300 tokens) and *vote* call (T=0.3, <=200 tokens) through
`VLLMAgent.batch_generate_json`, with the reference's retry schedule (main.py:256-478: up to 3
batched attempts per phase; with random-init weights every attempt runs to the token limit and
fails validation — the "retry storm" of SURVEY.md F7).  Prompts are the (system, user, schema)
triples the unmodified reference builds in steady-state rounds 6.. (tests/golden/
bench_prompts_*.json.gz, produced by oracle/tools/make_bench_prompts.py); each step uses the next
recorded round, so only what really repeats in a game (per-agent system prompts, identical retry
prompts) hits the prefix cache.

Printed JSON (one line, rank 0):
  value   agent-steps/s over device time only (CUDA-event time of prefill + decode; inputs
          already tokenised and resident)
  e2e     agent-steps/s wall-clock through VLLMAgent.batch_generate_json from host strings to
          parsed dicts (tokenisation, H2D, D2H, detokenise, json parse inside the timed region)
  roofline   the dominant kernel (decode-path tcgen05 GEMM streaming the gate_up weights), timed
             live with CUDA events over all 40 layers' weights (1.4e10 B > L2)
  decode  whole-step decode roofline (weights + KV bytes per step / step time)
  cpu_baseline  the CPU oracle (transformers fp32 + xgrammar) timed on this host, bounded sample
                (`extrapolated: true`: a 2-layer slice scaled to the full depth and retry schedule)
  fixed   the same metric on a FIXED schedule (every call of a round runs all rows to max_tokens,
          forced tokens "{" + blanks — the retry storm with nothing ever validating): identical work
          in every run, on every rank and in every round, so it can be compared round over round
          and its 1->N curve measures the machine rather than each game's retry luck
  ranks   per-rank device / wall seconds of the faithful region (value uses the max, as the
          contract says; value_mean_over_ranks is beside it)
  single_game (N > 1)  ONE game whose rows are sharded over the N replicas (dp.run_sharded): the
          latency half of SURVEY.md §8e; `value` above is the aggregate half (one game per replica)
"""
from __future__ import annotations

import argparse
import gzip
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

HF_NAME = {"qwen3-14b": "Qwen/Qwen3-14B", "qwen3-8b": "Qwen/Qwen3-8B", "qwen3-0.6b": "Qwen/Qwen3-0.6B",
           "qwen3-32b": "Qwen/Qwen3-32B", "mistral-22b": "mistralai/Mistral-Small-Instruct-2409"}
LLM = {"temperature_decide": 0.5, "temperature_vote": 0.3, "max_tokens_decide": 300,
       "max_tokens_vote": 200}          # reference config.py:52-58
MAX_RETRIES = 3                          # main.py:269, 386
BATCH_RETRY_THRESHOLD = 0.3              # main.py:270, 387


def load_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops", 1590.0), "measured"
    return 6650.0, 1590.0, "fallback"


def load_sustained_tf():
    """cuBLAS bf16 back to back for seconds (power-capped clocks) — the figure that applies to a GEMM
    timed inside a long step; None when the driver has not measured it."""
    p = ROOT / "MEASURED_PEAKS.json"
    return json.loads(p.read_text()).get("bf16_tflops_sustained") if p.exists() else None


def load_rounds(honest: int, byz: int, family: str = "qwen3"):
    p = ROOT / "tests" / "golden" / f"bench_prompts_{family}_{honest}h{byz}b.json.gz"
    with gzip.open(p, "rt") as f:
        return json.load(f)["rounds"]


# ---- host-side mirror of the reference's callers (main.py:232-254, 256-478) -----------------
def valid_decision(r) -> bool:
    if r is None or "error" in r or r.get("value") is None:
        return False
    a, b = r.get("internal_strategy", ""), r.get("public_reasoning", "")
    return isinstance(a, str) and len(a.strip()) >= 3 and isinstance(b, str) and len(b.strip()) >= 10


def valid_vote(r) -> bool:
    return r is not None and "error" not in r and r.get("decision", "") in ("stop", "continue")


def run_phase(agent, prompts, temperature, max_tokens, validator, counters):
    pending = list(range(len(prompts)))
    for attempt in range(1, MAX_RETRIES + 1):
        if not pending:
            break
        res = agent.batch_generate_json([tuple(prompts[i]) for i in pending], temperature=temperature,
                                        max_tokens=max_tokens)
        counters["calls"] += 1
        counters["rows"] += len(pending)
        failed = [i for i, r in zip(pending, res) if not validator(r)]
        counters["ok"] += len(pending) - len(failed)
        pending = failed
        if pending and attempt < MAX_RETRIES and len(pending) / len(prompts) <= BATCH_RETRY_THRESHOLD:
            # the reference switches to per-agent sequential retries here (up to 3 single calls
            # each, bcg_agents.py:683-791); the "RETRY ATTEMPT" prompt suffix is not replayed
            for i in list(pending):
                s, u, sc = prompts[i]
                for _ in range(MAX_RETRIES):
                    r = agent.generate_json(u, sc, temperature, max_tokens, system_prompt=s)
                    counters["calls"] += 1
                    counters["rows"] += 1
                    if validator(r):
                        counters["ok"] += 1
                        pending.remove(i)
                        break
            break
    counters["abstain"] += len(pending)


def agent_step(agent, rnd, counters):
    run_phase(agent, rnd["decide"]["prompts"], LLM["temperature_decide"], LLM["max_tokens_decide"],
              valid_decision, counters)
    run_phase(agent, rnd["vote"]["prompts"], LLM["temperature_vote"], LLM["max_tokens_vote"],
              valid_vote, counters)


def fixed_round(agent, engine, rnd, acc=None):
    """One game round on the FIXED schedule: MAX_RETRIES batched attempts per phase, every row
    teacher-forced through "{" + blanks (valid in all four reference grammars, never accepting),
    so every call runs all its rows to max_tokens.  Returns the number of engine calls."""
    from byzantine_consensus_llm_agents_b200.engine import Request
    tk = engine.tk
    brace, blank = tk.encode("{", False), tk.encode(" ", False)
    assert len(brace) == 1 and len(blank) == 1, "tokenizer must have single tokens for '{' and ' '"
    calls = 0
    for phase, temp, mt in (("decide", LLM["temperature_decide"], LLM["max_tokens_decide"]),
                            ("vote", LLM["temperature_vote"], LLM["max_tokens_vote"])):
        forced = brace + blank * (mt - 1)
        for _ in range(MAX_RETRIES):
            reqs = [Request(agent._format_chat_prompt(s, u), sc, temp, mt, forced=forced)
                    for s, u, sc in rnd[phase]["prompts"]]
            s0 = dict(engine.stats)
            res = (agent.llm.tp.generate if agent.llm.tp is not None else engine.generate)(reqs)
            if acc is not None:      # decode time per phase (the two phases differ in context length)
                acc.setdefault(phase, [0.0, 0])
                acc[phase][0] += engine.stats["decode_s"] - s0["decode_s"]
                acc[phase][1] += engine.stats["decode_steps"] - s0["decode_steps"]
            short = [(len(r.token_ids), r.finish) for r in res if len(r.token_ids) != mt]
            if short:
                raise RuntimeError(f"fixed schedule: {len(short)} of {len(res)} {phase} rows did not run to {mt} "
                                   f"tokens: {short[:4]}")
            calls += 1
    return calls


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "200", "-i", str(gpu_index)], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons, power = [], [], set(), []
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
                power.append(float(c[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        busy = [s for s, p in zip(sm, power) if p > 300] or sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---- dominant-kernel roofline, timed live ----------------------------------------------------
def kernel_roofline(engine, B: int, hbm_peak: float, peak_kind: str):
    import torch
    from byzantine_consensus_llm_agents_b200 import lib as L
    cfg, w = engine.cfg, engine.w
    N, K = 2 * cfg.d_ff, cfg.d_model
    x = engine.buf["xn"][:B]
    out = engine.buf["act"][:B]
    splits = L.splits_for(B, N, K)
    epi = L.EPI_SWIGLU if splits == 1 else L.EPI_F32
    if splits > 1:
        out = engine.buf["partials"][:splits * B * N].view(splits, B, N)
    layers = len(w.w_gate_up)
    # launched the way the decode step launches it: programmatic dependent launch when the engine
    # uses it (a launch's weight prefetch then overlaps its predecessor's drain, as in the step)
    gflags = engine.flags & L.GEMM_PDL
    for l in range(min(3, layers)):
        L.gemm(x, w.w_gate_up[l], out, epi, splits=splits, flags=gflags)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 3
    e0.record()
    for _ in range(reps):
        for l in range(layers):          # 40 different weight matrices: 14 GB >> 126 MB L2
            L.gemm(x, w.w_gate_up[l], out, epi, splits=splits, flags=gflags)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / (reps * layers)
    alg_bytes = N * K * 2 + B * K * 2 + (B * (N // 2) * 2 if splits == 1 else splits * B * N * 4)
    achieved = alg_bytes / (ms * 1e-3) / 1e9
    traffic, traffic_source = None, None
    prof = ROOT / "profiles" / "r02_gemm_decode_traffic.json"
    if prof.exists():
        try:
            pj = json.loads(prof.read_text())
            # the capture is of the 14B shape (weights dominate: rows <= 16 share the tile width)
            if (N, K) == tuple(pj.get("N_K", (34816, 5120))) and B <= 16:
                traffic = pj.get("dram_bytes_per_launch")
            traffic_source = ("profiles/r02_gemm_decode_traffic.json (one ncu --set full capture of this kernel "
                              "and shape, dram__bytes_read.sum + dram__bytes_write.sum; not re-measured in this run)")
        except Exception:
            traffic = None
        if traffic is None:
            traffic_source = None
    return {"kernel": f"gemm_tcgen05_kernel<BN={16 if B <= 16 else 32 if B <= 32 else 64 if B <= 64 else 128}> "
                      f"(decode gate_up: M={B}, N={N}, K={K}, splits={splits}"
                      f"{', back-to-back launches with programmatic dependent launch' if gflags else ''})",
            "bound": "hbm", "achieved": round(achieved, 1), "peak": hbm_peak, "peak_kind": peak_kind,
            "unit": "GB/s", "frac": round(achieved / hbm_peak, 4), "traffic": traffic,
            "traffic_source": traffic_source,
            "us_per_launch": round(ms * 1e3, 2), "algorithmic_bytes_per_launch": alg_bytes}


# ---- CPU baseline (oracle port) --------------------------------------------------------------
class CpuOracleSampler:
    """The CPU oracle (transformers fp32 forward with KV cache; xgrammar + sampler cost is included in
    the decode steps) on a BOUNDED sample of the workload: a `layers_sample`-layer, full-width model,
    one `prefill_tokens`-token prefill + `decode_steps` guided decode steps at B = 1 per sample
    (exactly how the oracle engine — and, with mixed schemas, the reference — runs a row).  Each
    sample is extrapolated to one full agent step: per-layer and head costs are separated, the
    layer cost is scaled to the full depth, and the retry schedule 3 x (decide prompt + 300) +
    3 x (vote prompt + 200) per agent is applied."""

    def __init__(self, model_key: str, rounds, layers_sample: int = 2, decode_steps: int = 48,
                 prefill_tokens: int = 1024):
        import torch
        from byzantine_consensus_llm_agents_b200.model import PRESETS, ModelConfig
        from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer
        from oracle.engine import OracleEngine
        from oracle.model import build_hf_model
        self.full = full = PRESETS[model_key]
        self.layers_sample, self.decode_steps, self.prefill_tokens = layers_sample, decode_steps, prefill_tokens
        cfg = ModelConfig(**{**full.__dict__, "n_layers": layers_sample, "name": full.name + "-sample"})
        self.cfg = cfg
        tk = get_tokenizer(cfg.family, cfg.vocab)
        # torchrun exports OMP_NUM_THREADS=1; the CPU arm is meant to use every host core
        try:
            torch.set_num_threads(os.cpu_count() or 1)
        except Exception:
            pass
        self.threads = torch.get_num_threads()
        self.hf = build_hf_model(cfg, seed=0)
        self.eng = OracleEngine(self.hf, tk, base_seed=0, memoise=False)
        s, u, self.schema = rounds[0]["decide"]["prompts"][0]
        if full.family == "mistral":
            prompt = f"<s>[INST] <<SYS>>\n{s}\n<</SYS>>\n\n{u} [/INST]"
        else:
            prompt = f"<|im_start|>system\n{s}<|im_end|>\n<|im_start|>user\n{u} /no_think<|im_end|>\n<|im_start|>assistant\n"
        ids = tk.encode(prompt)
        self.n_prompt = len(ids)
        self.short = tk.decode(ids[:prefill_tokens], skip_special_tokens=False)
        v = rounds[0]["vote"]["prompts"][0]
        self.n_vote = len(tk.encode(v[1])) + len(tk.encode(v[0]))
        with torch.no_grad():   # the head (final norm + lm_head) separately from the layers
            h = torch.randn(1, 1, cfg.d_model)
            t0 = time.perf_counter()
            for _ in range(3):
                self.hf.lm_head(self.hf.model.norm(h))
            self.t_head = (time.perf_counter() - t0) / 3

    def sample(self):
        """One bounded sample -> (wall seconds, extrapolated agent-steps/s, extrapolated decode tok/s)."""
        from oracle.ref_harness import GuidedDecodingParams, SamplingParams
        st0 = dict(self.eng.stats)
        sp = SamplingParams(temperature=0.5, max_tokens=self.decode_steps,
                            guided_decoding=GuidedDecodingParams(json=self.schema))
        t0 = time.perf_counter()
        self.eng([self.short], sp)
        t_sample = time.perf_counter() - t0
        st = {k: self.eng.stats[k] - st0[k] for k in st0}
        t_prefill_tok = st["prefill_s"] / max(1, st["prefill_tokens"])
        t_step = st["decode_s"] / max(1, st["decode_steps"])
        t_host_per_tok = max(0.0, (t_sample - st["prefill_s"] - st["decode_s"]) / self.decode_steps)
        t_layer_step = max(0.0, (t_step - self.t_head) / self.layers_sample)
        step_full = self.t_head + self.full.n_layers * t_layer_step + t_host_per_tok
        prefill_full_tok = t_prefill_tok * self.full.n_layers / self.layers_sample   # layers dominate prefill
        per_agent = MAX_RETRIES * (self.n_prompt * prefill_full_tok + LLM["max_tokens_decide"] * step_full) + \
            MAX_RETRIES * (self.n_vote * prefill_full_tok + LLM["max_tokens_vote"] * step_full)
        return t_sample, 1.0 / per_agent, 1.0 / step_full

    def describe(self, t_sample: float) -> str:
        return (f"oracle engine (transformers fp32 + xgrammar + oracle sampler), {self.layers_sample}-layer "
                f"full-width {self.full.name}: {self.prefill_tokens}-token prefill + {self.decode_steps} guided "
                f"decode steps at B=1 ({t_sample:.1f}s per sample), extrapolated to {self.full.n_layers} layers and "
                f"to 3x(decide {self.n_prompt} tok + 300) + 3x(vote {self.n_vote} tok + 200) per agent")


def cpu_baseline(model_key: str, rounds, n_agents: int, samples: int = 1, warmup: int = 0):
    """`samples` bounded samples of the CPU oracle (after `warmup` untimed ones); value = median of
    the per-sample extrapolations."""
    cs = CpuOracleSampler(model_key, rounds)
    for _ in range(warmup):
        cs.sample()
    got = [cs.sample() for _ in range(max(1, samples))]
    walls = [g[0] for g in got]
    return {"value": round(statistics.median(g[1] for g in got), 6), "unit": "agent-steps/s", "cores": cs.threads,
            "host_cpus": os.cpu_count(), "kind": "port", "extrapolated": True,
            "samples": len(got), "measured_sample_s": round(statistics.mean(walls), 2),
            "decode_tok_s": round(statistics.median(g[2] for g in got), 3),
            "sample": cs.describe(statistics.mean(walls))}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="qwen3-14b")
    ap.add_argument("--honest", type=int, default=8)
    ap.add_argument("--byzantine", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--fixed-rounds", type=int, default=1,
                    help="extra rounds on the fixed (all rows to max_tokens) schedule, reported under 'fixed'")
    ap.add_argument("--single-game-steps", type=int, default=1,
                    help="N > 1: extra rounds of ONE game with rows sharded over the replicas ('single_game')")
    ap.add_argument("--tp", type=int, default=1,
                    help="tensor_parallel_size of each replica.  Plain `python bench.py --gpus T --tp T`: the "
                         "adapter spawns ranks 1..T-1 (like vLLM's mp executor).  Under torchrun with N ranks: "
                         "N/T replicas of T consecutive ranks each, one game per replica (TP composed with DP)")
    ap.add_argument("--fused-rope", type=int, default=0,
                    help="1: qk-norm / RoPE / KV write inside decode attention (one launch fewer per layer)")
    ap.add_argument("--pdl", type=int, default=int(os.environ.get("BCG_B200_PDL", "1")),
                    help="programmatic dependent launch for the decode step (1 = on)")
    a = ap.parse_args()

    if os.environ.get("BCG_BENCH_WATCHDOG_S"):
        # a hung collective must not eat a GPU lease: dump every thread's stack and exit
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["BCG_BENCH_WATCHDOG_S"]), exit=True)
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    n_agents = a.honest + a.byzantine
    rounds = load_rounds(a.honest, a.byzantine, "mistral" if a.model.startswith("mistral") else "qwen3")
    hbm_peak, tf_peak, peak_kind = load_peaks()
    config = {"workload": f"{a.model} bf16 random-init, {a.honest} honest + {a.byzantine} Byzantine agents, "
                          "one game round per step (decide T=0.5/300 tok + vote T=0.3/200 tok, reference "
                          "retry schedule), recorded steady-state reference prompts",
              "agents": n_agents, "rounds_recorded": len(rounds), "pdl": bool(a.pdl),
              "parallelism": (f"tp{a.tp} x dp{max(1, world // a.tp)} (one game per tensor-parallel replica; decode "
                              "all-reduces fused into the norm-statistics kernels over NVLink peer memory)"
                              if a.tp > 1 else
                              f"dp{world} (one independent game per GPU)" if world > 1 else "single GPU"),
              "l2": "per-step inputs (all layer weights, tens of GB per decode step) exceed the 126 MB L2"}

    if a.impl == "reference":
        if rank != 0:
            return
        # The reference's own engine (vLLM) cannot run here; its CPU
        # path is the oracle port, timed on the host cores.
        # One STEP of this arm = one bounded sample (about 3-4 s of CPU work: 1,024-token prefill + 48
        # guided decode steps on a 2-layer full-width slice); W warm-up samples, then exactly K timed
        # ones.  `value` is the median of the per-sample extrapolations to a full agent step
        # (cpu_baseline.extrapolated = true); ms_per_step is the measured time of a sample.
        t0 = time.perf_counter()
        cb = cpu_baseline(a.model, rounds, n_agents, samples=a.steps, warmup=a.warmup)
        line = {"impl": "reference", "metric": "agent_steps_per_sec", "value": cb["value"],
                "unit": "agent-steps/s", "n_gpus": 0, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": round(1e3 * cb["measured_sample_s"], 1), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config, "cpu_baseline": cb,
                "full_step_ms_extrapolated": round(1e3 * n_agents / cb["value"], 1),
                "e2e": {"value": cb["value"], "unit": "agent-steps/s", "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0},
                "wall_s": round(time.perf_counter() - t0, 1)}
        print(json.dumps(line), flush=True)
        return

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    if world > 1:
        # No device_id when tensor-parallel replicas are cut out of the world: with an eagerly bound
        # NCCL communicator torch makes every NON-member of a later new_group() take part in an
        # ncclCommSplit that the members of a gloo group never join (a hang).
        kw = {} if a.tp > 1 else {"device_id": torch.device(f"cuda:{local_rank}")}
        dist.init_process_group("nccl", **kw)

    from byzantine_consensus_llm_agents_b200 import vllm_agent as VA
    model_name = HF_NAME.get(a.model, a.model)
    model_config = {"max_model_len": 8192, "gpu_memory_utilization": 0.9, "tensor_parallel_size": 1,
                    "max_num_seqs": 4, "disable_qwen3_thinking": True, "device": f"cuda:{local_rank}",
                    "seed": 1000 + rank, "engine_max_rows": max(16, n_agents), "pdl": bool(a.pdl),
                    "fused_rope": bool(a.fused_rope)}
    # collectives of the measurement itself: the default NCCL group, or — when tensor-parallel
    # replicas are cut out of the torchrun world — a gloo group of the replica leaders
    leaders, lg, red_dev = world, None, "cuda"
    if a.tp > 1:
        model_config["tensor_parallel_size"] = a.tp
        if world > 1:
            from byzantine_consensus_llm_agents_b200 import tp as TP
            dist.barrier()                                           # NCCL communicator over all ranks
            group, pos, replica, leader = TP.replica_groups(a.tp)
            leaders = world // a.tp
            lg = dist.new_group(list(range(0, world, a.tp)), backend="gloo")
            red_dev = "cpu"
            model_config.update(tp_group=group, tp_leader=leader, device=f"cuda:{local_rank - pos}",
                                seed=1000 + replica)
            if pos > 0:
                TP.serve_worker(model_name, model_config, pos, a.tp)   # returns when the leader closes
                dist.barrier()
                dist.destroy_process_group()
                return

    def barrier():
        if world > 1:
            dist.barrier(group=lg)

    t_load = time.perf_counter()
    agent = VA.VLLMAgent("agent_0", model_name, model_config)
    engine = agent.llm.engine
    torch.cuda.synchronize()
    t_load = time.perf_counter() - t_load

    counters = dict(calls=0, rows=0, ok=0, abstain=0)
    for i in range(a.warmup):
        agent_step(agent, rounds[i % len(rounds)], counters)
    torch.cuda.synchronize()
    barrier()
    counters = dict(calls=0, rows=0, ok=0, abstain=0)
    s0 = dict(engine.stats)
    clocks = ClockSampler(local_rank) if rank == 0 else None
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(a.steps):
        agent_step(agent, rounds[(a.warmup + i) % len(rounds)], counters)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    clk = clocks.stop() if clocks else None
    s1 = dict(engine.stats)
    d = {k: s1[k] - s0[k] for k in s1}
    dev_s = d["prefill_s"] + d["decode_s"]

    t = torch.tensor([wall, dev_s], dtype=torch.float64, device=red_dev)
    per_rank = [t.clone() for _ in range(leaders)]
    if world > 1:
        barrier()
        dist.all_gather(per_rank, t, group=lg)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=lg)
    wall_max, dev_max = t.tolist()
    total_agent_steps = n_agents * a.steps * leaders
    rank_wall = [round(float(x[0]), 3) for x in per_rank]
    rank_dev = [round(float(x[1]), 3) for x in per_rank]

    # ---- the fixed schedule (identical work everywhere; see the module docstring)
    fixed = None
    fixed_error = None
    fixed_phases = {}
    if a.fixed_rounds > 0:
        if a.tp == 1:
            engine.reset_prefix_cache()
        f0 = f1 = dict(engine.stats)
        try:
            fixed_round(agent, engine, rounds[0])                   # warm-up (graphs of this shape)
            torch.cuda.synchronize()
            barrier()
            f0 = dict(engine.stats)
            for i in range(a.fixed_rounds):
                fixed_round(agent, engine, rounds[(1 + i) % len(rounds)], fixed_phases)
            torch.cuda.synchronize()
            f1 = dict(engine.stats)
        except RuntimeError as e:        # reported in the line, never a reason to lose the faithful number
            fixed_error = str(e)
            if world > 1:
                raise
        fd = {k: f1[k] - f0[k] for k in f1}
        ft = torch.tensor([fd["prefill_s"] + fd["decode_s"]], dtype=torch.float64, device=red_dev)
        if world > 1:
            dist.all_reduce(ft, op=dist.ReduceOp.MAX, group=lg)
        fixed = {"error": fixed_error} if fixed_error else {
                 "value": round(n_agents * a.fixed_rounds * leaders / float(ft[0]), 4), "unit": "agent-steps/s",
                 "rounds": a.fixed_rounds, "decode_steps_per_round": fd["decode_steps"] / a.fixed_rounds,
                 "rows_per_step": round(fd["generated_tokens"] / max(1, fd["decode_steps"]), 2),
                 "decode_ms_per_step": round(1e3 * fd["decode_s"] / max(1, fd["decode_steps"]), 3),
                 "prefill_s_per_round": round(fd["prefill_s"] / a.fixed_rounds, 3),
                 "decode_ms_per_step_by_phase": {k: round(1e3 * v[0] / max(1, v[1]), 3)
                                                 for k, v in fixed_phases.items()},
                 "schedule": f"{MAX_RETRIES} batched attempts per phase, all {n_agents} rows forced to max_tokens"}

    # ---- ONE game, rows sharded over the replicas (latency half of SURVEY.md §8e)
    single = None
    if world > 1 and a.tp == 1 and a.single_game_steps > 0:
        engine.reset_prefix_cache()
        engine.base_seed, engine.calls = 4242, 0                    # every replica plays the SAME game
        agent.llm.data_parallel = True
        cs = dict(calls=0, rows=0, ok=0, abstain=0)
        agent_step(agent, rounds[0], cs)                            # warm-up (graphs of the sharded shapes)
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        for i in range(a.single_game_steps):
            agent_step(agent, rounds[(1 + i) % len(rounds)], cs)
        torch.cuda.synchronize()
        st = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        dist.all_reduce(st, op=dist.ReduceOp.MAX)
        agent.llm.data_parallel = False
        single = {"value": round(n_agents * a.single_game_steps / float(st[0]), 4), "unit": "agent-steps/s",
                  "steps": a.single_game_steps, "ms_per_step": round(1e3 * float(st[0]) / a.single_game_steps, 1),
                  "what": f"one game, its rows split round-robin over {world} replicas (no device collective; "
                          "texts gathered host-side); compare with e2e / n_gpus"}

    if rank == 0:
        cfg = engine.cfg
        gen, steps = d["generated_tokens"], d["decode_steps"]
        # whole-step decode roofline: W + sum_rows S*KVB per step
        avg_rows = d["generated_tokens"] / max(1, d["decode_steps"])
        avg_ctx = (d["prompt_tokens"] / max(1, d["prompts"])) + 0.5 * (gen / max(1, d["prompts"]))
        # per GPU: under TP every rank streams 1/tp of the weights and of the KV heads
        bytes_per_step = (cfg.decode_weight_bytes() + avg_rows * avg_ctx * cfg.kv_bytes_per_token) / a.tp
        step_s = d["decode_s"] / max(1, steps)
        dec_ach = bytes_per_step / step_s / 1e9 if steps else 0.0
        pf_flops = d["prefill_tokens"] * cfg.prefill_gemm_flops_per_token() / a.tp   # per GPU
        tf_sus = load_sustained_tf()
        line = {
            "metric": "agent_steps_per_sec", "value": round(total_agent_steps / dev_max, 4),
            "unit": "agent-steps/s", "n_gpus": world if world > 1 else a.tp, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": round(1e3 * wall_max / a.steps, 2), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": config,
            "e2e": {"value": round(total_agent_steps / wall_max, 4), "unit": "agent-steps/s",
                    "h2d_bytes_per_step": int(d["h2d_bytes"] / a.steps),
                    "d2h_bytes_per_step": int(d["d2h_bytes"] / a.steps)},
            "gpu_launches": int(d["gpu_launches"]),
            "decode_tok_s": round(leaders * gen / max(1e-9, d["decode_s"]), 1),
            "decode": {"bound": "hbm", "bytes_per_step": int(bytes_per_step), "ms_per_step": round(step_s * 1e3, 3),
                       "rows_per_step": round(avg_rows, 2), "avg_context": int(avg_ctx),
                       "achieved": round(dec_ach, 1), "peak": hbm_peak, "unit": "GB/s",
                       "frac": round(dec_ach / hbm_peak, 4), "frac_of_8TBs": round(dec_ach / 8000.0, 4)},
            "prefill": {"bound": "tensor", "new_tokens": int(d["prefill_tokens"]),
                        "cached_tokens": int(d["cached_tokens"]), "seconds": round(d["prefill_s"], 3),
                        "achieved": round(pf_flops / max(1e-9, d["prefill_s"]) / 1e12, 1), "peak": tf_peak,
                        "unit": "TFLOP/s (GEMM flops / whole prefill time incl. attention, norms, sampling)",
                        "frac": round(pf_flops / max(1e-9, d["prefill_s"]) / 1e12 / tf_peak, 4),
                        "peak_kind": "burst (best of 10 cuBLAS launches)", "peak_sustained": tf_sus,
                        "frac_of_sustained": (round(pf_flops / max(1e-9, d["prefill_s"]) / 1e12 / tf_sus, 4)
                                              if tf_sus else None)},
            "value_mean_over_ranks": round(n_agents * a.steps * leaders * leaders / max(1e-9, sum(rank_dev)), 4),
            "ranks": {"device_s": rank_dev, "wall_s": rank_wall,
                      "note": "one independently seeded game per rank; value divides by the slowest rank"},
            "fixed": fixed, "single_game": single,
            "per_step": {"engine_calls": counters["calls"] / a.steps, "rows": counters["rows"] / a.steps,
                         "valid": counters["ok"] / a.steps, "abstain": counters["abstain"] / a.steps,
                         "generated_tokens": gen / a.steps, "decode_steps": steps / a.steps,
                         "host_ms": round(1e3 * (wall - dev_s) / a.steps, 1)},
            "model_load_s": round(t_load, 1), "clocks": clk,
        }
        if a.tp == 1:
            line["roofline"] = kernel_roofline(engine, n_agents, hbm_peak, peak_kind)
        if world == 1 and a.tp == 1 and not a.no_cpu_baseline:
            VA.VLLMAgent.shutdown()
            try:
                line["cpu_baseline"] = cpu_baseline(a.model, rounds, n_agents)
            except Exception as e:  # the baseline is reported, never a reason to lose the GPU number
                line["cpu_baseline"] = {"value": None, "unit": "agent-steps/s", "cores": os.cpu_count(),
                                        "kind": "port", "sample": f"failed: {e}"}
        print(json.dumps(line), flush=True)
    if a.tp > 1:
        VA.VLLMAgent.shutdown()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
