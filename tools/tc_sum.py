#!/usr/bin/env python
"""One line per engine of tools/tc_report.py's stdout (stdin): accuracy against the oracle and timing."""
import json, sys
for l in sys.stdin:
    if " {" not in l:
        continue
    n, j = l.split(" ", 1)
    d = json.loads(j)
    print(n, "max %.2e ever-open %.2e median %.2e above %d audio %.2e sqlmis %d  kernel %.1f us step %.1f us" % (
        d["ch_iq_rel_rms_max"], d["ch_iq_rel_rms_max_channels_ever_open"], d["ch_iq_rel_rms_median"], d["channel_blocks_above_1e-4"],
        d["audio_rel_rms_max"], d["squelch_decisions_differing"], d["channelizer_kernel_us"], d["us_per_block"]))
