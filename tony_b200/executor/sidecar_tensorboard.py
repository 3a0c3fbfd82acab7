#!/usr/bin/env python
"""Sidecar TensorBoard task: serve ``TB_LOG_DIR`` on the reserved ``TB_PORT``.

Counterpart of tony-core/src/main/resources/sidecar_tensorboard.py:20-31.  If the
``tensorboard`` package is absent (it is not on the B200 image) a plain HTTP file server over
the log directory keeps the tracking URL alive; ``SIDECAR_TB_TEST`` makes it exit at once.
"""
import os
import sys


def main() -> int:
    log_dir = os.environ.get("TB_LOG_DIR", ".")
    port = int(os.environ.get("TB_PORT", "6006"))
    if os.environ.get("SIDECAR_TB_TEST"):
        print(f"sidecar tensorboard test mode: log_dir={log_dir} port={port}")
        return 0
    try:
        from tensorboard import program  # type: ignore

        tb = program.TensorBoard()
        tb.configure(argv=[None, "--logdir", log_dir, "--port", str(port), "--bind_all"])
        tb.main()
        return 0
    except ImportError:
        import functools
        import http.server

        os.makedirs(log_dir, exist_ok=True)
        handler = functools.partial(http.server.SimpleHTTPRequestHandler, directory=log_dir)
        http.server.ThreadingHTTPServer(("0.0.0.0", port), handler).serve_forever()
        return 0


if __name__ == "__main__":
    sys.exit(main())
