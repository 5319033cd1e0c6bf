#!/bin/bash
python tools/small_batch_probe.py 2>&1 | tail -6
