// Game rules for TicTacToe, Connect4 and DragonChess over packed bitboards (sm_100a).
//
// Each game is a struct with the same static interface so the rule kernels (rules.cu), the
// evaluators (eval.cu, net.cu) and the tree search (search.cu) are written once:
//
//   State                     packed position (4 / 16 / 40 bytes), also holds Player/PreviousPlayer
//   A, H, W, C                action count and network-input shape (planes are [1,H,W,C] int8)
//   initial()                 start position
//   player(s), prev(s)        side to move (1|2) and PreviousPlayer (0 = None)
//   legal_row(s, from) -> u64 legal targets grouped by "row": action ids are enumerated in
//                             ascending order as (row ascending, bit ascending); NROWS rows
//   action_of(row, bit)       action id;  row_of(a)/bit_of(a) inverse
//   apply(s, a) -> bool       false = illegal ("Tried to make an illegal move.")
//   winner(s, prev_action)    -1 None, 0 draw (python int 0), 1|2 win, 3 = np.float64(0.0)
//                             (the reference's empty-cell artefact)
//   plane(s, cell, ch)        network input value at [0, cell/W, cell%W, ch]
//
// Reference behaviour restated here (file:line in /root/reference/src):
//   TicTacToe.py:29-36 (legal) :41-48 (apply) :50-55 (planes) :57-105 (winner)
//   Connect4.py:30-36 :41-53 :55-60 :62-108
//   DragonChess.py:78-106 :111-125 :127-214 :161-167 :242-371
#pragma once
#include <stdint.h>

#ifndef __CUDACC__
#define __host__
#define __device__
#endif
#define BB_HD __host__ __device__ __forceinline__

namespace bb {

BB_HD int popc64(uint64_t v) {
#ifdef __CUDA_ARCH__
  return __popcll(v);
#else
  return __builtin_popcountll(v);
#endif
}
BB_HD int ctz64(uint64_t v) {   // v != 0
#ifdef __CUDA_ARCH__
  return __ffsll((long long)v) - 1;
#else
  return __builtin_ctzll(v);
#endif
}

// ------------------------------------------------------------------------------------------
// TicTacToe: bits 0-8 X (plane 0), 9-17 O (plane 1), bit 18 = Player-1, bits 19-20 = PreviousPlayer
// ------------------------------------------------------------------------------------------
struct TicTacToe {
  struct State { uint32_t v; };
  static constexpr int A = 9, H = 3, W = 3, C = 3, NROWS = 1, ID = 0;

  BB_HD static State initial() { return State{0u}; }
  BB_HD static int player(const State& s) { return 1 + ((s.v >> 18) & 1); }
  BB_HD static int prev(const State& s) { return (s.v >> 19) & 3; }
  BB_HD static void set_meta(State& s, int player, int prev) {
    s.v = (s.v & 0x3FFFFu) | ((uint32_t)(player - 1) << 18) | ((uint32_t)prev << 19);
  }
  BB_HD static int next_player(const State& s) { return 3 - player(s); }
  BB_HD static uint32_t xs(const State& s) { return s.v & 0x1FF; }
  BB_HD static uint32_t os(const State& s) { return (s.v >> 9) & 0x1FF; }

  BB_HD static uint64_t legal_row(const State& s, int) { return (~(xs(s) | os(s))) & 0x1FF; }
  BB_HD static int action_of(int, int bit) { return bit; }
  BB_HD static int row_of(int) { return 0; }
  BB_HD static int bit_of(int a) { return a; }

  BB_HD static bool apply(State& s, int a) {
    if (a < 0 || a >= A || (((xs(s) | os(s)) >> a) & 1)) return false;
    int p = player(s);
    s.v |= 1u << (a + (p == 1 ? 0 : 9));
    set_meta(s, p == 2 ? 1 : 2, p);
    return true;
  }

  // cells lying on a full line of `b`
  BB_HD static uint32_t on_line(uint32_t b) {
    const uint32_t L[8] = {0x007, 0x038, 0x1C0, 0x049, 0x092, 0x124, 0x111, 0x054};
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r |= ((b & L[i]) == L[i]) ? L[i] : 0u;
    return r;
  }

  BB_HD static int winner(const State& s, int prev_action) {
    uint32_t x = xs(s), o = os(s);
    uint32_t wx = on_line(x), wo = on_line(o);
    if (prev_action >= 0) {
      uint32_t bit = 1u << prev_action;
      if (x & bit) { if (wx & bit) return 1; }
      else if (o & bit) { if (wo & bit) return 2; }
      else if (on_line(~(x | o) & 0x1FF) & bit) return 3;       // run of EMPTY cells: float 0.0
    } else {
      uint32_t w = wx | wo;                                      // row-major first winning cell
      if (w) return ((wx >> ctz64(w)) & 1) ? 1 : 2;
    }
    return ((x | o) == 0x1FF) ? 0 : -1;
  }

  BB_HD static int plane(const State& s, int cell, int ch) {
    if (ch == 0) return (xs(s) >> cell) & 1;
    if (ch == 1) return (os(s) >> cell) & 1;
    return player(s) == 1 ? 1 : -1;
  }
};

// ------------------------------------------------------------------------------------------
// Connect4: x/o bitboards, bit = col*7 + row (row 0 = bottom, bit 6 of each column always 0).
// Meta lives in o: bit 56 = Player-1, bits 57-58 = PreviousPlayer.
// ------------------------------------------------------------------------------------------
struct Connect4 {
  struct __align__(16) State { uint64_t x, o; };
  static constexpr int A = 7, H = 6, W = 7, C = 3, NROWS = 1, ID = 1;
  static constexpr uint64_t BOARD = 0x0000FDFBF7EFDFBFull;   // 6 rows x 7 cols
  static constexpr uint64_t TOPROW = 0x0000810204081020ull;  // row 5 of every column
  static constexpr uint64_t ROW0 = 0x0000040810204081ull;

  BB_HD static State initial() { return State{0ull, 0ull}; }
  BB_HD static int player(const State& s) { return 1 + (int)((s.o >> 56) & 1); }
  BB_HD static int prev(const State& s) { return (int)((s.o >> 57) & 3); }
  BB_HD static void set_meta(State& s, int player, int prev) {
    s.o = (s.o & BOARD) | ((uint64_t)(player - 1) << 56) | ((uint64_t)prev << 57);
  }
  BB_HD static int next_player(const State& s) { return 3 - player(s); }
  BB_HD static uint64_t xs(const State& s) { return s.x & BOARD; }
  BB_HD static uint64_t os(const State& s) { return s.o & BOARD; }

  BB_HD static uint64_t legal_row(const State& s, int) {
    uint64_t top = ~(xs(s) | os(s)) & TOPROW;                   // Connect4.py:33: top cell empty
    uint64_t m = 0;
#pragma unroll
    for (int j = 0; j < 7; ++j) m |= ((top >> (j * 7 + 5)) & 1ull) << j;
    return m;
  }
  BB_HD static int action_of(int, int bit) { return bit; }
  BB_HD static int row_of(int) { return 0; }
  BB_HD static int bit_of(int a) { return a; }

  // highest occupied row of a column, -1 if empty (works for non-gravity bit patterns too)
  BB_HD static int top_of(uint64_t occ, int col) {
    uint32_t c = (uint32_t)(occ >> (col * 7)) & 0x3F;
    int t = -1;
#pragma unroll
    for (int i = 0; i < 6; ++i) if ((c >> i) & 1) t = i;
    return t;
  }

  BB_HD static bool apply(State& s, int a) {
    if (a < 0 || a >= A) return false;
    uint64_t occ = xs(s) | os(s);
    int t = top_of(occ, a);
    if (t == 5) return false;                                    // Connect4.py:42-43
    int p = player(s);
    uint64_t bit = 1ull << (a * 7 + t + 1);
    if (p == 1) s.x |= bit; else s.o |= bit;
    set_meta(s, p == 2 ? 1 : 2, p);
    return true;
  }

  // cells lying on a run of >= 4 of `b` (vertical 1, horizontal 7, diagonals 8 and 6)
  BB_HD static uint64_t on_line(uint64_t b) {
    uint64_t r = 0;
    const int D[4] = {1, 7, 6, 8};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int d = D[i];
      uint64_t m = b & (b >> d);
      m = m & (m >> (2 * d));                                    // run of 4 starts here
      r |= m | (m << d) | (m << (2 * d)) | (m << (3 * d));
    }
    return r;
  }

  BB_HD static int winner(const State& s, int prev_action) {
    uint64_t x = xs(s), o = os(s), occ = x | o;
    if (prev_action >= 0) {
      int t = top_of(occ, prev_action);
      if (t < 0) return 3;              // empty column: six empty cells in a vertical line (A.5)
      uint64_t bit = 1ull << (prev_action * 7 + t);
      if (x & bit) { if (on_line(x) & bit) return 1; }
      else if (on_line(o) & bit) return 2;
    } else {
      uint64_t wx = on_line(x), wo = on_line(o), w = wx | wo;
      if (w) {                          // row-major scan order: rows ascending, columns ascending
#pragma unroll
        for (int i = 0; i < 6; ++i) {
          uint64_t r = w & (ROW0 << i);
          if (r) return ((wx >> ctz64(r)) & 1) ? 1 : 2;
        }
      }
    }
    return ((occ & TOPROW) == TOPROW) ? 0 : -1;                  // _isOver: top row full
  }

  BB_HD static int plane(const State& s, int cell, int ch) {
    int bit = (cell % 7) * 7 + cell / 7;
    if (ch == 0) return (int)((s.x >> bit) & 1);
    if (ch == 1) return (int)((s.o >> bit) & 1);
    return player(s) == 1 ? 1 : -1;
  }
};

// ------------------------------------------------------------------------------------------
// DragonChess: 8 rows x 8 nibbles (0 empty, 1-6 white K P N B R Q, 9-14 black), row 0 = White's
// home rank.  meta: bits 0-3 castle WK WQ BK BQ, bit 4 = Player-1, bits 5-6 = PreviousPlayer.
// Actions: ordered pairs of distinct squares, id = sq1*63 + sq2 - (sq2 > sq1).
// ------------------------------------------------------------------------------------------
struct DragonChess {
  struct __align__(8) State { uint32_t row[8]; uint32_t meta; uint32_t pad; };
  static constexpr int A = 4032, H = 8, W = 8, C = 17, NROWS = 64, ID = 2;
  enum { K = 1, P = 2, N = 3, B = 4, R = 5, Q = 6, BLACK = 8 };

  BB_HD static State initial() {
    State s;
    s.row[0] = 0x00010000u;                       // K on (0,4)
    s.row[1] = 0x00222000u;                       // P on (1,3..5)
    s.row[2] = s.row[3] = s.row[4] = s.row[5] = 0u;
    s.row[6] = 0xAAAAAAAAu;                       // black pawns
    s.row[7] = 0xDBC9ECBDu;                       // r n b q k b n r  (col 0 = low nibble)
    s.meta = 0xCu;                                // BK, BQ castle flags; Player 1; Prev None
    s.pad = 0u;
    return s;
  }
  BB_HD static int player(const State& s) { return 1 + ((s.meta >> 4) & 1); }
  BB_HD static int prev(const State& s) { return (s.meta >> 5) & 3; }
  BB_HD static int next_player(const State& s) { return (prev(s) == 1 && player(s) == 1) ? 2 : 1; }
  BB_HD static int at(const State& s, int r, int c) { return (s.row[r] >> (4 * c)) & 0xF; }
  BB_HD static void put(State& s, int r, int c, int v) {
    s.row[r] = (s.row[r] & ~(0xFu << (4 * c))) | ((uint32_t)v << (4 * c));
  }
  BB_HD static bool is_white(int v) { return v >= 1 && v <= 6; }
  BB_HD static bool is_black(int v) { return v >= 9; }

  // legal targets (bit = row*8+col) of the piece on square `from`
  BB_HD static uint64_t legal_row(const State& s, int from) {
    int pc = at(s, from >> 3, from & 7);
    if (pc == 0) return 0;
    bool white = is_white(pc);
    if (white != (player(s) == 1)) return 0;     // _sanity_check: wrong side to move
    uint64_t own = 0, foe = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      uint32_t w = s.row[r];
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        int v = (w >> (4 * c)) & 0xF;
        if (v) {
          if (is_white(v) == white) own |= 1ull << (r * 8 + c); else foe |= 1ull << (r * 8 + c);
        }
      }
    }
    return legal_row_occ(s, from, own, foe);
  }
#ifdef __CUDACC__
  // The same with the occupancy of the side to move (own) and of the other side (foe) worked out ONCE per position by the
  // whole warp (lane l looks at squares l and l + 32, two ballots per colour) instead of by a 64-step loop in every call:
  // an expansion asks for 64 rows, and the loop was most of its instructions.
  static constexpr bool HAS_OCC = true;
  __device__ static void occ_warp(const State& s, uint64_t* own, uint64_t* foe) {
    const int l = threadIdx.x & 31;
    const int v0 = at(s, l >> 3, l & 7), v1 = at(s, 4 + (l >> 3), l & 7);
    const uint64_t wh = (uint64_t)__ballot_sync(0xFFFFFFFFu, is_white(v0)) | ((uint64_t)__ballot_sync(0xFFFFFFFFu, is_white(v1)) << 32);
    const uint64_t bl = (uint64_t)__ballot_sync(0xFFFFFFFFu, is_black(v0)) | ((uint64_t)__ballot_sync(0xFFFFFFFFu, is_black(v1)) << 32);
    const bool w = player(s) == 1;
    *own = w ? wh : bl;
    *foe = w ? bl : wh;
  }
  __device__ static int winner_warp(const State& s, int) {    // winner(): both kings looked for by the warp
    const int l = threadIdx.x & 31;
    const int v0 = at(s, l >> 3, l & 7), v1 = at(s, 4 + (l >> 3), l & 7);
    const bool wk = __any_sync(0xFFFFFFFFu, v0 == K || v1 == K), bk = __any_sync(0xFFFFFFFFu, v0 == (BLACK | K) || v1 == (BLACK | K));
    return !bk ? 1 : (!wk ? 2 : -1);
  }
#endif
  // own / foe: occupancy of the colour of the piece on `from` / of the other colour; 0 when the piece is not the mover's
  BB_HD static uint64_t legal_row_occ(const State& s, int from, uint64_t own, uint64_t foe) {
    int r1 = from >> 3, c1 = from & 7;
    int pc = at(s, r1, c1);
    if (pc == 0) return 0;
    bool white = is_white(pc);
    if (white != (player(s) == 1)) return 0;     // _sanity_check: wrong side to move
    int kind = pc & 7;
    uint64_t occ = own | foe, out = 0;
    const uint64_t bit = 1ull << from;
    constexpr uint64_t FA = 0x0101010101010101ull, FH = 0x8080808080808080ull;    // columns 0 and 7
    if (kind == K) {
      // one step in the eight directions: shifts of the piece's bit, wraps across the board edge masked at the destination
      out = (bit << 8) | (bit >> 8) | (((bit << 1) | (bit << 9) | (bit >> 7)) & ~FA) | (((bit >> 1) | (bit << 7) | (bit >> 9)) & ~FH);
      return out & ~own;                          // castling shapes are never legal (A.19)
    }
    if (kind == N) {
      out = (((bit << 17) | (bit >> 15)) & ~FA) | (((bit << 15) | (bit >> 17)) & ~FH) | (((bit << 10) | (bit >> 6)) & ~(FA | (FA << 1))) |
            (((bit << 6) | (bit >> 10)) & ~(FH | (FH >> 1)));
      return out & ~own;
    }
    if (kind == P) {
      int fwd = white ? 1 : -1, home = white ? 1 : 6, last = white ? 6 : 1;
      if (r1 == last) return 0;                   // promotion shapes return 0 => frozen pawn
      int r2 = r1 + fwd;
      if (r2 < 0 || r2 > 7) return 0;
      if (r1 == home) {
        if (!((occ >> (r2 * 8 + c1)) & 1)) {
          out |= 1ull << (r2 * 8 + c1);
          int r3 = r2 + fwd;
          if (!((occ >> (r3 * 8 + c1)) & 1)) out |= 1ull << (r3 * 8 + c1);
        }
      } else if (r1 > 1 && r1 < 6) {
        if (!((occ >> (r2 * 8 + c1)) & 1)) out |= 1ull << (r2 * 8 + c1);
      }
      if (c1 > 0 && ((foe >> (r2 * 8 + c1 - 1)) & 1)) out |= 1ull << (r2 * 8 + c1 - 1);
      if (c1 < 7 && ((foe >> (r2 * 8 + c1 + 1)) & 1)) out |= 1ull << (r2 * 8 + c1 + 1);
      return out;
    }
    // sliders: along each ray up to and including the first blocker, then without the own pieces -- as occluded fills
    // (Kogge-Stone, three doubling steps per direction: no loop over squares, no branch per step)
    const uint64_t empty = ~occ;
    const uint64_t gr = kind != B ? bit : 0ull, gb = kind != R ? bit : 0ull;      // rook-wise / bishop-wise movers
    out = ray_up<8>(gr, empty, ~0ull) | ray_down<8>(gr, empty, ~0ull) | ray_up<1>(gr, empty, ~FA) | ray_down<1>(gr, empty, ~FH) |
          ray_up<9>(gb, empty, ~FA) | ray_up<7>(gb, empty, ~FH) | ray_down<7>(gb, empty, ~FA) | ray_down<9>(gb, empty, ~FH);
    return out & ~own;
  }
  // squares attacked from `gen` along one direction (shift by S towards higher / lower bits), stopping at the first non-empty
  // square (included); m = squares a step may land on (the column a wrap would land on is excluded)
  template <int S>
  BB_HD static uint64_t ray_up(uint64_t gen, uint64_t pro, uint64_t m) {
    pro &= m;
    gen |= pro & (gen << S);
    pro &= pro << S;
    gen |= pro & (gen << (2 * S));
    pro &= pro << (2 * S);
    gen |= pro & (gen << (4 * S));
    return (gen << S) & m;
  }
  template <int S>
  BB_HD static uint64_t ray_down(uint64_t gen, uint64_t pro, uint64_t m) {
    pro &= m;
    gen |= pro & (gen >> S);
    pro &= pro >> S;
    gen |= pro & (gen >> (2 * S));
    pro &= pro >> (2 * S);
    gen |= pro & (gen >> (4 * S));
    return (gen >> S) & m;
  }
  BB_HD static int action_of(int from, int to) { return from * 63 + to - (to > from ? 1 : 0); }
  BB_HD static int row_of(int a) { return a / 63; }
  BB_HD static int bit_of(int a) { int f = a / 63, r = a - f * 63; return r + (r >= f ? 1 : 0); }

  BB_HD static bool apply(State& s, int a) {
    if (a < 0 || a >= A) return false;            // ids >= 4032 (promotion/castle) are never legal
    int from = row_of(a), to = bit_of(a);
    if (!((legal_row(s, from) >> to) & 1ull)) return false;
    apply_trusted(s, a);
    return true;
  }
  // the move itself, for an action that is known to be legal (a child of the tree: it was generated from legal_row)
  BB_HD static void apply_trusted(State& s, int a) {
    int from = row_of(a), to = bit_of(a);
    int pc = at(s, from >> 3, from & 7);
    put(s, to >> 3, to & 7, pc);
    put(s, from >> 3, from & 7, 0);
    int pl = player(s), pv = prev(s);
    if (pv == 1 && pl == 1) { pl = 2; pv = 1; } else { pv = pl; pl = 1; }   // DragonChess.py:192-197
    uint32_t f = s.meta & 0xF;                    // :199-212
    if (at(s, 0, 4) != K) f &= ~3u; else if (at(s, 0, 7) != R) f &= ~1u;
    if (at(s, 0, 0) != R) f &= ~2u;
    if (at(s, 7, 4) != (BLACK | K)) f &= ~12u; else if (at(s, 7, 7) != (BLACK | R)) f &= ~4u;
    if (at(s, 7, 0) != (BLACK | R)) f &= ~8u;
    s.meta = f | ((uint32_t)(pl - 1) << 4) | ((uint32_t)pv << 5);
  }

  BB_HD static int winner(const State& s, int) {
    bool wk = false, bk = false;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      uint32_t w = s.row[r];
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        int v = (w >> (4 * c)) & 0xF;
        wk |= (v == K);
        bk |= (v == (BLACK | K));
      }
    }
    return !bk ? 1 : (!wk ? 2 : -1);              // DragonChess.py:161-167
  }

  BB_HD static int plane(const State& s, int cell, int ch) {
    if (ch < 12) {
      int v = at(s, cell >> 3, cell & 7);
      if (v == 0) return 0;
      // piece_map (DragonChess.py:11): P0 p1 R2 r3 N4 n5 B6 b7 Q8 q9 K10 k11
      const int base[7] = {0, 10, 0, 4, 6, 2, 8};
      return ch == base[v & 7] + (v >> 3);
    }
    if (ch < 16) return (s.meta >> (ch - 12)) & 1;
    return (player(s) == 1 && prev(s) == 1) ? 1 : 0;
  }
};

// Sentinel winner codes shared with the host shims.
enum { WIN_NONE = -1, WIN_DRAW = 0, WIN_P1 = 1, WIN_P2 = 2, WIN_FLOAT0 = 3 };

}  // namespace bb
