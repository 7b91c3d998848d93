"""Writes tests/golden/delegator_session.json: the log of oracle/delegator_script.py run against the
UNMODIFIED reference TaskDelegator (/root/reference).  Test infrastructure.
usage (container with /root/reference): python oracle/gen_golden_delegator.py"""
import asyncio
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")
from pilott.delegation.task_delegator import TaskDelegator          # noqa: E402
from oracle.delegator_script import session                          # noqa: E402


def main():
    log = asyncio.run(session(TaskDelegator))
    path = os.path.join(ROOT, "tests", "golden", "delegator_session.json")
    with open(path, "w") as f:
        json.dump(log, f, indent=1)
    print(path, len(log), "entries")


if __name__ == "__main__":
    main()
