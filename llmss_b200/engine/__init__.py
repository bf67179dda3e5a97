"""Inference engine: paged KV cache, ModelRunner (prefill + CUDA-graph decode), scheduler."""
