"""Drop-in check against the reference's own classes, where the reference tree is present (this
container; skipped on the GPU box, which has no /root/reference).  For every class this build
mirrors: same public method names, same positional parameters in the same order with the same
defaults (extras must be keyword-only), same pydantic fields, defaults and bounds.  Host-only:
nothing here touches the GPU or the CUDA library."""
import inspect
import os
import sys

import pytest

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "pilott")), reason="reference tree not present")


@pytest.fixture(scope="module")
def ref():
    sys.path.insert(0, REF)
    try:
        import pilott.memory.enhanced_memory as mem
        import pilott.knowledge.knowledge_manager as km
        import pilott.core.router as router
        import pilott.orchestration.load_balancer as lb
        import pilott.delegation.task_delegator as dg
        yield {"mem": mem, "km": km, "router": router, "lb": lb, "dg": dg}
    finally:
        sys.path.remove(REF)


def _positional(fn):
    ps = [p for p in inspect.signature(fn).parameters.values()
          if p.kind in (p.POSITIONAL_ONLY, p.POSITIONAL_OR_KEYWORD)]
    return [(p.name, p.default) for p in ps]


def _check_methods(ours, theirs, names):
    for n in names:
        assert hasattr(ours, n), f"{ours.__name__}.{n} missing"
        got, want = _positional(getattr(ours, n)), _positional(getattr(theirs, n))
        assert got[: len(want)] == want, f"{ours.__name__}.{n}: {got} vs reference {want}"
        extra = got[len(want):]
        assert not extra, f"{ours.__name__}.{n}: extra positional parameters {extra} (extras must be keyword-only)"
        assert inspect.iscoroutinefunction(getattr(ours, n)) == inspect.iscoroutinefunction(getattr(theirs, n)), n


def _check_model(ours, theirs):
    fo, ft = ours.model_fields, theirs.model_fields
    assert set(ft) <= set(fo), f"{ours.__name__}: missing fields {set(ft) - set(fo)}"
    for name, f in ft.items():
        g = fo[name]
        assert g.annotation == f.annotation or str(g.annotation) == str(f.annotation), (ours.__name__, name)
        if not f.is_required() and f.default_factory is None:
            assert g.default == f.default, (ours.__name__, name, g.default, f.default)
        assert sorted(map(repr, g.metadata)) == sorted(map(repr, f.metadata)), (ours.__name__, name, g.metadata, f.metadata)


def test_enhanced_memory_interface(ref):
    from pilottai_b200.memory import EnhancedMemory, MemoryItem
    them = ref["mem"]
    _check_methods(EnhancedMemory, them.EnhancedMemory,
                   ["__init__", "start", "stop", "store_semantic", "semantic_search", "store_task", "store_interaction",
                    "store_pattern", "get_pattern", "get_recent_tasks", "cleanup", "clear"])
    _check_model(MemoryItem, them.MemoryItem)


def test_knowledge_manager_interface(ref):
    from pilottai_b200.knowledge import CacheEntry, KnowledgeManager, KnowledgeSource
    them = ref["km"]
    _check_methods(KnowledgeManager, them.KnowledgeManager,
                   ["__init__", "start", "stop", "add_source", "query_knowledge", "invalidate_cache", "cleanup",
                    "get_source_stats", "_generate_cache_key", "_is_cache_entry_valid"])
    _check_model(CacheEntry, them.CacheEntry)
    _check_model(KnowledgeSource, them.KnowledgeSource)


def test_router_interface(ref):
    from pilottai_b200.orchestration import RouterConfig, TaskPriority, TaskRouter
    them = ref["router"]
    _check_methods(TaskRouter, them.TaskRouter, ["__init__", "route_task", "get_task_priority"])
    _check_model(RouterConfig, them.RouterConfig)
    assert {p.name: p.value for p in TaskPriority} == {p.name: p.value for p in them.TaskPriority}


def test_load_balancer_interface(ref):
    from pilottai_b200.orchestration import LoadBalancer, LoadBalancerConfig, LoadMetrics
    them = ref["lb"]
    _check_methods(LoadBalancer, them.LoadBalancer,
                   ["__init__", "start", "stop", "get_metrics", "_find_best_agent", "_calculate_composite_load",
                    "_calculate_load_trend", "_analyze_agent_loads", "_redistribute_tasks", "_move_task",
                    "_get_moveable_tasks", "_is_task_moveable", "_update_metrics_history"])
    _check_model(LoadBalancerConfig, them.LoadBalancerConfig)
    _check_model(LoadMetrics, them.LoadMetrics)


def test_task_delegator_interface(ref):
    from pilottai_b200.orchestration import TaskDelegator
    them = ref["dg"]
    _check_methods(TaskDelegator, them.TaskDelegator,
                   ["__init__", "start", "stop", "evaluate_delegation", "record_delegation", "get_agent_metrics",
                    "get_metrics", "_find_best_agent", "_should_delegate", "_periodic_cleanup", "_cleanup_old_history"])


def test_every_public_method_of_the_reference_classes_exists_here(ref):
    from pilottai_b200.knowledge import KnowledgeManager
    from pilottai_b200.memory import EnhancedMemory
    from pilottai_b200.orchestration import LoadBalancer, TaskDelegator, TaskRouter
    pairs = [(EnhancedMemory, ref["mem"].EnhancedMemory), (KnowledgeManager, ref["km"].KnowledgeManager),
             (TaskRouter, ref["router"].TaskRouter), (LoadBalancer, ref["lb"].LoadBalancer),
             (TaskDelegator, ref["dg"].TaskDelegator)]
    for ours, theirs in pairs:
        public = [n for n, v in inspect.getmembers(theirs, callable) if not n.startswith("_")]
        missing = [n for n in public if not hasattr(ours, n)]
        assert not missing, f"{ours.__name__} lacks public methods of the reference: {missing}"
