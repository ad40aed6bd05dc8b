// libyb200_yafray.so — the one exported symbol exporters look up in the reference's interface plugin:
// `extern "C" yafrayInterface_t* getYafray()` (src/interface/yafrayinterface.cc:386-392,
// typedef interfaceConstructor include/interface/yafrayinterface.h:148), here returning the C++ host class of
// include/yb200_interface.hpp over the C ABI of libyb200.so.
#define YB200_DEFINE_GETYAFRAY
#include "../../include/yb200_interface.hpp"
