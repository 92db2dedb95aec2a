from .gpt import GPT, GPTConfig, flatten_parameters

__all__ = ["GPT", "GPTConfig", "flatten_parameters"]
