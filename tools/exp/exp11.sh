#!/bin/bash
L=gpurun_out/exp11.log
: > $L
timeout 300 build/attn_harness quick 2>&1 | grep -E "FAIL|SUMMARY|error|WATCHDOG" | cut -c1-220 >> $L
timeout 300 build/attn_harness bwd 2>&1 | grep -E "FAIL|SUMMARY|error|WATCHDOG" | cut -c1-220 >> $L
timeout 300 build/attn_harness multi 2>&1 | grep -E "FAIL|SUMMARY|error|WATCHDOG" | cut -c1-220 >> $L
timeout 120 build/attn_harness perfall >> $L 2>&1
