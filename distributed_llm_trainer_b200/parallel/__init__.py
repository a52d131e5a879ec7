"""Parallel engines: flat parameter storage (``flat``), NVLink comm runtime (``comm``), data
parallelism (``ddp``) and fully-sharded data parallelism (``fsdp``)."""
from .flat import ALIGN, FlatParamStore, plan_layout

__all__ = ["ALIGN", "FlatParamStore", "plan_layout"]
