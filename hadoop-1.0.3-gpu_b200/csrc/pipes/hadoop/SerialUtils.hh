// Source-compatibility shim: Pipes programs include "hadoop/SerialUtils.hh".
#include "../hgm/pipes_api.hh"
