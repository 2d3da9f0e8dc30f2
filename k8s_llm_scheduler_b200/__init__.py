"""k8s_llm_scheduler_b200 -- B200-native fallback placement path of K8s-LLM-Scheduler.

One hot path only (reference scheduler.py:521-559 and the table arithmetic that feeds it):
filter -> score -> first-best argmax, as hand-written sm_100a CUDA behind a C ABI
(include/k8s_place_b200.h).  This package holds the CUDA sources (csrc/), their build
recipe, the ctypes binding and the host-side mirror of the reference interface.

There is no CPU fallback: importing the ops without the built CUDA library raises.
"""
from ._lib import (FILTER_CAPACITY, FILTER_READY, STRATEGY_LEAST_LOADED,
                   STRATEGY_RESOURCE_BALANCED, STRATEGY_ROUND_ROBIN, PlacementError)
from .tables import NodeMetrics, NodeTable, PodSpec, PodTable, SchedulingDecision, strategy_id

__all__ = ["FILTER_CAPACITY", "FILTER_READY", "STRATEGY_LEAST_LOADED", "STRATEGY_RESOURCE_BALANCED",
           "STRATEGY_ROUND_ROBIN", "PlacementError", "NodeMetrics", "NodeTable", "PodSpec",
           "PodTable", "SchedulingDecision", "strategy_id"]
