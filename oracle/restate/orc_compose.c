/* oracle/restate/orc_compose.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 * Restates the sort-last composite: DirectSendBase::composeLayered
 * (DirectSend/Base/DirectSendBase.cpp:316-458), the direct-send round
 * (:76-281), getPieceRange (:59-74), the depth-swap 'over'
 * (Common/ImageRGBAFloatColorDepthSort.hpp:13-27) and Gather
 * (Common/ImageColorOnly.hpp:220-270), evaluated pixel by pixel. */
#include "oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

void orc_piece_range(int num_pixels, int piece, int num_pieces, int* begin,
                     int* end) {
  const int size = num_pixels / num_pieces;
  *begin = size * piece;
  *end = (piece < num_pieces - 1) ? (*begin + size) : num_pixels;
}

void orc_blend(const float* top, const float* bottom, float* out) {
  const float td = top[4], bd = bottom[4];
  const int top_front = td <= bd;
  const float* front = top_front ? top : bottom;
  const float* back = top_front ? bottom : top;
  float r[5];
  for (int c = 0; c < 4; ++c) r[c] = front[c] + back[c] * (1.0f - front[3]);
  r[4] = (bd < td) ? bd : td; /* std::min(topDepth, bottomDepth) */
  memcpy(out, r, sizeof(r));
}

typedef struct {
  float depth;
  int owner, local, layer;
} entry_t;

static int entry_cmp(const void* pa, const void* pb) { /* :380-388 */
  const entry_t* a = (const entry_t*)pa;
  const entry_t* b = (const entry_t*)pb;
  if (a->depth == b->depth) {
    if (a->owner == b->owner) return (a->local > b->local) - (a->local < b->local);
    return (a->owner > b->owner) - (a->owner < b->owner);
  }
  return (a->depth < b->depth) ? -1 : 1;
}

void orc_compose(int nranks, int width, int height, int nlayers,
                 const int* layer_owner, const float* const* layer_data,
                 const float* depth_hints, const int* rank_order, float* full5,
                 int* piece_owner) {
  const int npix = width * height;
  static const float kClear[5] = {0.0f, 0.0f, 0.0f, 0.0f, INFINITY};

  /* :363-388 global (depth, owner, localIndex) order */
  entry_t* order = (entry_t*)malloc(sizeof(entry_t) * (size_t)(nlayers + 1));
  int* next_local = (int*)calloc((size_t)nranks, sizeof(int));
  int n = 0;
  for (int proc = 0; proc < nranks; ++proc)
    for (int l = 0; l < nlayers; ++l)
      if (layer_owner[l] == proc) {
        order[n].depth = depth_hints[l];
        order[n].owner = proc;
        order[n].local = next_local[proc]++;
        order[n].layer = l;
        ++n;
      }
  qsort(order, (size_t)n, sizeof(entry_t), entry_cmp);

  /* position of each world rank in the ordered group */
  int* position = (int*)malloc(sizeof(int) * (size_t)nranks);
  for (int i = 0; i < nranks; ++i) position[rank_order[i]] = i;

  for (int piece = 0; piece < nranks; ++piece) {
    int begin, end;
    orc_piece_range(npix, piece, nranks, &begin, &end);
    for (int p = begin; p < end; ++p) {
      if (piece_owner) piece_owner[p] = rank_order[piece];
      float acc[5];
      int have = 0;
      int idx = 0;
      while (idx < n) { /* :400-446 one direct-send round per same-owner run */
        const int owner = order[idx].owner;
        float run[5];
        memcpy(run, layer_data[order[idx].layer] + 5 * (size_t)p, sizeof(run));
        ++idx;
        while (idx < n && order[idx].owner == owner) { /* :413-426 */
          orc_blend(run, layer_data[order[idx].layer] + 5 * (size_t)p, run);
          ++idx;
        }
        /* :257-281: the receiver blends the pieces of all senders in group
         * order; every sender but the owner contributes a cleared frame */
        float round[5];
        for (int s = 0; s < nranks; ++s) {
          const float* img = (s == position[owner]) ? run : kClear;
          if (s == 0) {
            memcpy(round, img, sizeof(round));
          } else {
            orc_blend(round, img, round);
          }
        }
        if (!have) { /* :441-445 */
          memcpy(acc, round, sizeof(acc));
          have = 1;
        } else {
          orc_blend(acc, round, acc);
        }
      }
      if (!have) memcpy(acc, kClear, sizeof(acc)); /* :450-455 */
      memcpy(full5 + 5 * (size_t)p, acc, sizeof(acc));
    }
  }
  free(position);
  free(next_local);
  free(order);
}
