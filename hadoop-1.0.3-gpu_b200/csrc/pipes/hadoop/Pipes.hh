// Source-compatibility shim: Pipes programs include "hadoop/Pipes.hh".
#include "../hgm/pipes_api.hh"
