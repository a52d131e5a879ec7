from .dummy import create_dummy_dataloader
from .text import (OpenWebTextConfig, OpenWebTextDataset, OpenWebTextIterableDataset, TextDataConfig,
                   TinyStoriesConfig, TinyStoriesDataset, TinyStoriesIterableDataset,
                   create_openwebtext_dataloader, create_text_dataloader, create_tinystories_dataloader)
from .tokenizer import ByteTokenizer, load_tokenizer

__all__ = [n for n in dir() if not n.startswith("_")]
