"""Command-line entry points: ``run_sim`` (one simulation) and ``execute`` (batch driver)."""
