"""Multi-GPU communication over NVLink peer memory (see :mod:`llmtrain.parallel.nvl`)."""
