// msda_small.cuh -- placeholder: single-phase kernels for launch-latency-bound problems (filled in below)
#pragma once
