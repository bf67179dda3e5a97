"""llmss_b200 -- Blackwell (B200, sm_100a) native tensor-parallel LLM serving.

Capability parity target: jongwon-jay-lee/llmss (reference `src/llmss/`), re-designed
B200-first.  Layer map:

    serving/   broker (RESP LPUSH/RPOP/LLEN), HTTP producer, consumer, gRPC, scheduler
    engine/    paged KV cache, ModelRunner (prefill + CUDA-graph decode), sampler
    models/    gptj, gpt_bigcode (reference parity) + gpt2, opt, llama
    parallel/  TP shard plan, symmetric-memory communicator (P2P over NVLink)
    ops/       ctypes binding to the hand-written sm_100a kernels in csrc/ (+ CPU oracles)
    utils/     dist / hub / weights / layers  (the reference's public utility API)
"""

__version__ = "0.1.0"

from .models import MODEL_REGISTRY  # noqa: E402,F401
