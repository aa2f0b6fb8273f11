#!/bin/bash
python tools/e2e_probe.py 1; python tools/e2e_probe.py 2; python tools/e2e_probe.py 3
