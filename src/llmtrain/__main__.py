"""``python -m llmtrain`` entry (reference: `src/llmtrain/__main__.py:1-4`)."""

import sys

from llmtrain.cli import main

if __name__ == "__main__":
    sys.exit(main())
