#!/bin/bash
B=scripts/dev/scan1_bench
$B 100000 128 10 1.0 l2
$B 1000000 768 100 1.0 cosine
$B 10000000 384 10 0.8 l2
$B 10000000 384 10 0.01 l2
$B 4096 768 32 1.0 dot
$B 1000 64 10 1.0 l2
$B 300000 100 7 0.5 cosine
