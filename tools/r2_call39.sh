#!/bin/bash
timeout 120 python tools/lincomb_timing.py
