// oracle/shim/bctree/timer.hpp -- TEST INFRASTRUCTURE ONLY: the reference's timer macros compiled out (its own #else branch).
#pragma once
#define TimerStart(name)
#define TimerStop(name)
#define TimerLog(name)
#define TimerStopAndLog(name)
