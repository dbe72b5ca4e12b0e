"""Entry point: parse the config, join the job, train (the reference's L6, ``R:690`` / ``R:792``)."""
from __future__ import annotations

import json
import sys

from .config import parse_cli
from .trainer import Trainer


def main(argv=None) -> int:
    cfg = parse_cli(argv)
    trainer = Trainer(cfg)
    try:
        last = trainer.fit()
        if trainer.ctx.is_coordinator:
            print(json.dumps({k: v for k, v in last.items()}, default=float))
    except Exception:
        from .ops import loader
        note = loader.last_device_error()          # which barrier wait timed out, if a kernel trapped
        if note:
            print("device error record:", note, file=sys.stderr)
        raise
    finally:
        trainer.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
