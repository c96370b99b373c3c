// Oracle build only: see bgzf.h in this directory.
#pragma once
