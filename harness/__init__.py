"""Benchmark / test harness around the engine: stages the reference's model code
(`examples/`, `MinkowskiEngine/modules/`) unchanged into git-ignored
`baseline/_ref/`, provides import shims for the third-party packages those
scripts import at module top but that are absent from this image, and builds
the synthetic workloads of SURVEY.md §8d.  Not part of the engine."""
