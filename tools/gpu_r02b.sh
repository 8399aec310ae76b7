#!/bin/bash
# A/B: refill threshold of wf_traverse and pool size under the class-sorted scheduling
tools/gpu_ab.sh r02b tools/ab_r02b.spec
