from . import embeddings  # noqa: F401
