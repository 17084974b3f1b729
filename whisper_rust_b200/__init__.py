"""B200-native replacement for the whisper-rs transcription hot path."""
from . import synth  # noqa: F401
from .api import (  # noqa: F401
    DtwModelPreset, DtwParameters, FullParams, SamplingStrategy, WhisperContext,
    WhisperContextParameters, WhisperError, WhisperState, load_library,
)
