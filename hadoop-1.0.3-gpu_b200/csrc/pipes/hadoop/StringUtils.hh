// Source-compatibility shim: Pipes programs include "hadoop/StringUtils.hh".
#include "../hgm/pipes_api.hh"
