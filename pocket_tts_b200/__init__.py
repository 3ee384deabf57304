"""pocket_tts_b200 — B200-native batched implementation of Pocket TTS's per-frame generation
path (FlowLM decode step, flow head, streaming Mimi decoder) behind a C ABI."""

# the reference pins its public surface to these two names (tests/test_python_api.py:8-10); the engine classes
# (GenerationEngine, ModelConfig) stay importable by name
__all__ = ["TTSModel", "export_model_state"]


def __getattr__(name):
    if name in ("GenerationEngine", "ModelConfig"):
        from . import engine

        return getattr(engine, name)
    if name in ("TTSModel", "export_model_state"):
        from . import tts_model

        return getattr(tts_model, name)
    raise AttributeError(name)
