"""B200-native engine for the per-round agent step of the Byzantine Consensus Game.

Drop-in for the reference's `vllm_agent.py` (see vllm_agent.py in this package); the hot path is
the sm_100a library built from csrc/ and bound in lib.py.  Import as

    import byzantine_consensus_llm_agents_b200 as bcg
"""
__all__ = ["lib", "model", "build"]
