#!/bin/bash
timeout -s KILL 600 python tools/engine_first_vs_second.py 2>&1 | tail -12
