"""Smallest end-to-end case for compute-sanitizer: append, fused search (pair kernel), exact
fallback, routing.  usage: compute-sanitizer --tool memcheck python tools/sanitize_case.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pilottai_b200 import synth
from pilottai_b200.store import SemanticStore
from pilottai_b200.routing import RoutingEngine

x = synth.corpus(3000, 128, seed=1, dup_frac=0.01)
prio, tag, exp = synth.metadata(3000, seed=2)
st = SemanticStore(128, 2048, 0)
st.append(x[:1500], prio[:1500], tag[:1500], exp[:1500]); st.append(x[1500:], prio[1500:], tag[1500:], exp[1500:])
q = synth.queries(40, 128, seed=3)
for kern in ("tc_pair", "tc", "generic"):
    st.set_kernel(kern)
    idx, sc = st.search(q, 10, min_priority=1, now_us=1_000_000)
    idx2, sc2 = st.search(q, 40)                      # exact path
print("search ok", idx[0, :3], idx2[0, :3])
st.compact(1_000_000)
eng = RoutingEngine(synth.agent_table(40, seed=5), 0)
a, s = eng.route(synth.task_table(200, seed=6), "router", load_threshold=0.8)
print("route ok", a[:5])
st.close()
