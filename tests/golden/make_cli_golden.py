"""Dump the reference parser's flags to tests/golden/reference_cli.json (run in the build container):

    python tests/golden/make_cli_golden.py

Imports /root/reference/task.py with `google.cloud.storage` stubbed (it is not installed) and
intercepts ArgumentParser.parse_args to walk parser._actions.  The committed JSON lets the CLI-parity test run
where /root/reference does not exist (the GPU box).
"""
import argparse
import importlib.util
import json
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))


def reference_actions(path="/root/reference/task.py"):
    for name in ("google", "google.cloud", "google.cloud.storage"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["google"].cloud = sys.modules["google.cloud"]
    sys.modules["google.cloud"].storage = sys.modules["google.cloud.storage"]
    spec = importlib.util.spec_from_file_location("reference_task", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    captured = {}

    class Stop(Exception):
        pass

    def fake_parse(self, *a, **k):
        captured["parser"] = self
        raise Stop()

    orig = argparse.ArgumentParser.parse_args
    argparse.ArgumentParser.parse_args = fake_parse
    env_backup = {k: os.environ.pop(k, None) for k in ("AIP_MODEL_DIR", "WORLD_SIZE")}
    try:
        mod.main()
    except Stop:
        pass
    finally:
        argparse.ArgumentParser.parse_args = orig
        for k, v in env_backup.items():
            if v is not None:
                os.environ[k] = v
    return describe(captured["parser"])


def describe(parser, only=None):
    out = []
    for a in parser._actions:
        if isinstance(a, argparse._HelpAction):
            continue
        if only is not None and a.dest not in only:
            continue
        out.append({"flags": list(a.option_strings), "dest": a.dest, "type": getattr(a.type, "__name__", None),
                    "default": a.default, "action": type(a).__name__, "nargs": a.nargs,
                    "has_choices": a.choices is not None})
    return out


if __name__ == "__main__":
    acts = reference_actions()
    with open(os.path.join(HERE, "reference_cli.json"), "w") as fh:
        json.dump(acts, fh, indent=1)
    print(len(acts), "flags written")
