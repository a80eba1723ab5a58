"""Plugin contract of the modality layer — host-side mirror of src/multimeditron/model/modalities/base.py.

Same public names, argument meaning and error behaviour as the reference (BaseModalityConfig :10-35,
BaseModalityProcessor :37-79, BaseModality :81-161, AutoModality :164-222), so code and tests written against
the reference read the same here.  The classes are plain Python (+ torch.nn.Module): the B200 path does not need
HF's PreTrainedModel machinery to run; `multimeditron_b200.integration` registers the same modality classes
with the reference's own AutoModality (and thereby HF AutoModel/AutoConfig) when the reference is installed.
"""
from __future__ import annotations

import abc
import copy
from typing import Any, Dict, Optional

import torch


class BaseModalityConfig:
    """hidden_size = LLM hidden the projector maps to; modality_type = key of the modality ('image')."""

    model_type: str = ""

    def __init__(self, hidden_size: int = 1024, modality_type: Optional[str] = None, **kwargs):
        self.modality_type = modality_type
        self.hidden_size = hidden_size
        for k, v in kwargs.items():
            setattr(self, k, v)

    def to_dict(self) -> Dict[str, Any]:
        d = {k: copy.deepcopy(v) for k, v in self.__dict__.items() if not k.startswith("_")}
        d["model_type"] = type(self).model_type
        return d

    @classmethod
    def from_dict(cls, config: Dict[str, Any], **kwargs):
        d = {k: v for k, v in {**config, **kwargs}.items() if k not in ("model_type", "modality_type")}
        return cls(**d)


class BaseModalityProcessor(abc.ABC):
    """CPU-side preprocessing of one modality record (runs in DataLoader workers)."""

    def __init__(self, config: BaseModalityConfig):
        self.config = config

    @abc.abstractmethod
    def process(self, modality: Dict[str, Any]) -> Dict[str, Any]:
        """Return a copy of `modality` with 'value' = tensor and 'num_embeddings' = int (base.py:56-67)."""
        raise NotImplementedError

    def __call__(self, modality: Dict[str, Any]) -> Dict[str, Any]:
        return self.process(modality)


class BaseModality(torch.nn.Module, abc.ABC):
    """forward(inputs: list[Tensor]) -> Tensor[n_items, num_embeddings, hidden_size]."""

    config_class: type = BaseModalityConfig
    preprocessor_class: type = None

    def __init__(self, config: BaseModalityConfig, dtype: torch.dtype = torch.bfloat16):
        super().__init__()
        self.config = config
        self.tokenizer = None
        self._dtype = dtype

    @property
    def dtype(self) -> torch.dtype:
        return self._dtype

    def get_config(self) -> BaseModalityConfig:
        return self.config

    @abc.abstractmethod
    def freeze_modality_embedder(self): ...

    @abc.abstractmethod
    def unfreeze_modality_embedder(self): ...

    @abc.abstractmethod
    def unfreeze_projection(self): ...

    def freeze_all(self):
        self.freeze_modality_embedder()
        for p in self.parameters():
            p.requires_grad = False

    def unfreeze_all(self):
        self.unfreeze_projection()
        self.unfreeze_modality_embedder()
        for p in self.parameters():
            p.requires_grad = True


class AutoModality:
    """Name -> modality class registry (reference: base.py:164-222)."""

    _registry: Dict[str, type] = {}

    def __init__(self):
        raise RuntimeError("AutoModality should not be instantiated directly. Please use the 'from_name' method.")

    @classmethod
    def register(cls, name: str):
        def decorator(klass):
            if not (isinstance(klass, type) and issubclass(klass, BaseModality)):
                raise ValueError(f"Class {getattr(klass, '__name__', klass)} must inherit from BaseModality to be registered.")
            if name in cls._registry:
                raise ValueError(f"Modality name '{name}' is already registered.")
            if getattr(klass, "preprocessor_class", None) is None:
                raise ValueError(f"Modality class '{klass.__name__}' must define a 'preprocessor_class' attribute.")
            cls._registry[name] = klass
            klass.config_class.model_type = name
            return klass
        return decorator

    @classmethod
    def _lookup(cls, name: str) -> type:
        if name not in cls._registry:
            raise ValueError(f"Modality name '{name}' is not registered. Available values are {list(cls._registry.keys())}")
        return cls._registry[name]

    @classmethod
    def from_config(cls, config: BaseModalityConfig, **kwargs) -> BaseModality:
        """What `AutoModel.from_config(modality_config)` does for a registered modality (model.py:271)."""
        return cls._lookup(type(config).model_type)(config, **kwargs)

    @classmethod
    def preprocessor_from_name(cls, name: str, *args, **kwargs) -> BaseModalityProcessor:
        return cls._lookup(name).preprocessor_class(*args, **kwargs)

    @classmethod
    def config_from_dict(cls, config: dict, **kwargs) -> BaseModalityConfig:
        assert "model_type" in config, "Config dictionary must contain a 'model_type' key."
        return cls._lookup(config["model_type"]).config_class.from_dict(config, **kwargs)
