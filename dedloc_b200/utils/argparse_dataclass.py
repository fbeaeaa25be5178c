"""Dataclass -> argparse bridge with HfArgumentParser's CLI conventions (`--flag value`, booleans accept
true/false or no value, `List[str]` takes several values), used by every trainer entry point so that the
reference's flag names keep working without depending on transformers' TrainingArguments (whose field set
drifted: SURVEY 5.6). Unknown flags are accepted and ignored with a warning."""
from __future__ import annotations

import argparse
import dataclasses
import logging
import sys
import typing
from typing import Any, List, Optional, Sequence, Tuple

logger = logging.getLogger(__name__)


def _str2bool(v):
    if isinstance(v, bool):
        return v
    if v.lower() in ("yes", "true", "t", "y", "1"):
        return True
    if v.lower() in ("no", "false", "f", "n", "0"):
        return False
    raise argparse.ArgumentTypeError(f"expected a boolean, got {v!r}")


class DataclassArgumentParser(argparse.ArgumentParser):
    def __init__(self, dataclass_types: Sequence[type], **kwargs):
        kwargs.setdefault("formatter_class", argparse.ArgumentDefaultsHelpFormatter)
        super().__init__(**kwargs)
        self.dataclass_types = list(dataclass_types)
        seen = set()
        for dt in self.dataclass_types:
            hints = typing.get_type_hints(dt)
            for f in dataclasses.fields(dt):
                if not f.init or f.name in seen:
                    continue
                seen.add(f.name)
                self._add_field(f, hints[f.name])

    def _add_field(self, f: dataclasses.Field, tp) -> None:
        kw: dict = {"help": f.metadata.get("help")}
        origin = typing.get_origin(tp)
        if origin is typing.Union:  # Optional[X]
            args = [a for a in typing.get_args(tp) if a is not type(None)]
            tp = args[0] if args else str
            origin = typing.get_origin(tp)
        has_default = f.default is not dataclasses.MISSING or f.default_factory is not dataclasses.MISSING
        default = f.default if f.default is not dataclasses.MISSING else (
            f.default_factory() if f.default_factory is not dataclasses.MISSING else None)
        if tp is bool:
            kw.update(type=_str2bool, nargs="?", const=True, default=default if has_default else False)
        elif origin in (list, List):
            inner = (typing.get_args(tp) or (str,))[0]
            kw.update(type=inner, nargs="*", default=default if has_default else [])
        else:
            kw.update(type=tp if tp in (int, float, str) else str, default=default)
            if not has_default:
                kw["required"] = True
        self.add_argument(f"--{f.name}", **kw)

    def parse_args_into_dataclasses(self, args: Optional[Sequence[str]] = None) -> Tuple[Any, ...]:
        ns, unknown = self.parse_known_args(args)
        if unknown:
            logger.warning("ignoring unknown command-line arguments: %s", unknown)
        out = []
        for dt in self.dataclass_types:
            names = {f.name for f in dataclasses.fields(dt) if f.init}
            vals = {k: v for k, v in vars(ns).items() if k in names}
            for k, v in list(vals.items()):
                # HF convention: a single comma-separated string for list flags is also accepted
                if isinstance(v, list) and len(v) == 1 and isinstance(v[0], str) and "," in v[0]:
                    vals[k] = [x for x in v[0].split(",") if x]
            out.append(dt(**vals))
        return tuple(out)
