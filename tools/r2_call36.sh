#!/bin/bash
timeout 300 python tools/step_profile.py 2>&1 | grep -E "pack_|device time"
