// Stand-in for the Windows-only EasyX <graphics.h> (main.cpp:13,539-542,585-598): display and
// BMP saving only, no arithmetic on the hot path.  Nothing here is ever executed: the driver
// never calls the reference's main() (renamed ref_main by -Dmain=ref_main).
#pragma once
typedef unsigned int DWORD;
typedef unsigned int COLORREF;
typedef unsigned char BYTE;
typedef unsigned short WORD;
#define _T(x) x
#define BGR(c) (c)
static inline void initgraph(int, int) {}
static inline DWORD* GetImageBuffer() { static DWORD px[1280 * 720]; return px; }
static inline void FlushBatchDraw() {}
static inline void saveimage(const char*) {}
static inline void closegraph() {}
