"""pocket_tts_b200 — B200-native batched implementation of Pocket TTS's per-frame generation
path (FlowLM decode step, flow head, streaming Mimi decoder) behind a C ABI."""

__all__ = ["GenerationEngine", "ModelConfig"]


def __getattr__(name):
    if name in __all__:
        from . import engine

        return getattr(engine, name)
    raise AttributeError(name)
