"""pilottai_b200 -- B200-native hot path for PilottAI (semantic memory + agent routing).

Host-side mirrors of the reference API (import these instead of the `pilott.*` originals):
    EnhancedMemory, MemoryItem                      (pilott/memory/enhanced_memory.py)
    TaskRouter, RouterConfig, TaskPriority          (pilott/core/router.py)
    LoadBalancer, LoadBalancerConfig, LoadMetrics   (pilott/orchestration/load_balancer.py)
    TaskDelegator                                   (pilott/delegation/task_delegator.py)
    KnowledgeManager, KnowledgeSource               (pilott/knowledge/knowledge_manager.py)
`serve_adapter.attach(serve)` wires the router and the memory into a reference `Serve`.
Lower level: SemanticStore (store.py), RoutingEngine (routing.py), ShardedSemanticStore
(sharded.py).  Everything runs through libpilott_b200.so (include/pilott_b200.h); there is no
CPU fallback.  Submodules are imported lazily so that `import pilottai_b200` works on a box
without the built library (build.py lives here too).
"""
__version__ = "0.1.0"

_LAZY = {
    "EnhancedMemory": "memory", "MemoryItem": "memory",
    "TaskRouter": "orchestration", "RouterConfig": "orchestration", "TaskPriority": "orchestration",
    "LoadBalancer": "orchestration", "LoadBalancerConfig": "orchestration", "LoadMetrics": "orchestration",
    "TaskDelegator": "orchestration",
    "KnowledgeManager": "knowledge", "KnowledgeSource": "knowledge", "SemanticMemorySource": "knowledge",
    "SemanticStore": "store", "RoutingEngine": "routing", "ShardedSemanticStore": "sharded",
}


def __getattr__(name):
    mod = _LAZY.get(name)
    if mod is None:
        raise AttributeError(name)
    import importlib
    return getattr(importlib.import_module(f".{mod}", __name__), name)
