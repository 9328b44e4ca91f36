// Source-compatibility shim: Pipes programs include "hadoop/TemplateFactory.hh".
#include "../hgm/pipes_api.hh"
