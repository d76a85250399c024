"""brotli_lib_b200 — B200-native Brotli encoder hot path behind brotli-lib's encode API."""
from .encode import BrotliEncoder, Context, EncoderMode, brotliEncode, default_context  # noqa: F401

__all__ = ["brotliEncode", "BrotliEncoder", "EncoderMode", "Context", "default_context"]
