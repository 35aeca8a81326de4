"""Timers, stat monitor, profiler hooks, NaN/Inf guard, field/param dump."""
