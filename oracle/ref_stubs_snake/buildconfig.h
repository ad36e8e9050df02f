// TEST INFRASTRUCTURE ONLY. Stand-in for the header cmake generates from /root/reference/Snake/buildconfig.h.in:
// a default CPU build (no SNAKE_CUDA, no SNAKE_MOBILE, no sanitizer switches).
#pragma once
