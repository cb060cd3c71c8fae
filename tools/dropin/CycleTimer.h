// Stand-in for the reference's CycleTimer.h (rdtsc based; its MHz branch divides by an
// uninitialised value, CycleTimer.h:149): only currentSeconds() is used by the driver.
#ifndef DPSVM_DROPIN_CYCLETIMER_H
#define DPSVM_DROPIN_CYCLETIMER_H
#include <chrono>
class CycleTimer {
public:
    static double currentSeconds()
    {
        return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
    }
};
#endif
