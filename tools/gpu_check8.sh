#!/usr/bin/env bash
timeout 900 python tools/ab_onepass.py 2>&1 | grep layers
