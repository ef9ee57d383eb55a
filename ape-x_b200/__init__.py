"""B200-native prioritized-replay hot path for Bing-Jing/Ape-X.

This directory is meant to be put FIRST on sys.path: its modules carry the reference's
module names (`memory`, `batchrecorder`, `batchrecoder_AQL`) so the reference drivers import
them unchanged.  See DESIGN.md / INTEGRATION.md at the repo root."""
