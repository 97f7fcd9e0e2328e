/* TEST INFRASTRUCTURE ONLY (oracle/). Not part of the product path.
 *
 * The reference's unrooted-Newick reader is a flex/bison grammar
 * (libpll parse_utree.y / lex_utree.l); neither tool exists on this image, so
 * the reference objects are linked against this hand-written stand-in that
 * exports the symbols the bison epilogue would have provided:
 *
 *   pll_utree_parse_newick, pll_utree_parse_newick_string  (parse_utree.y:447-526)
 *   pll_utree_reset_template_indices                       (parse_utree.y:297-340)
 *   pll_utree_wraptree                                     (parse_utree.y:385-445)
 *   pll_utree_destroy, pll_utree_graph_destroy             (parse_utree.y:37-138)
 *
 * Behaviour that matters for parity and is reproduced here:
 *   - grammar: "(" sub "," sub "," sub ")" [label] [":" len] ";" at the top,
 *     strictly binary "(" sub "," sub ")" [label] [":" len] below, tips are
 *     label [":" len]                                      (parse_utree.y:179-246)
 *   - an inner node is a ring of three records; record 0 faces the parent,
 *     records 1 and 2 face the first / second listed child; child branch
 *     lengths are mirrored onto the parent-side records     (parse_utree.y:205-230)
 *   - a missing length is stored as 0                       (parse_utree.y:226,236)
 *   - tips are numbered 0.. in left-to-right order, inner nodes in post-order
 *     after them, the top-level trifurcation last           (parse_utree.y:250-340)
 */
#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "pll.h"

typedef struct
{
  const char * p;
  unsigned int tips;
  int failed;
} nwk_cursor;

static void nwk_fail(nwk_cursor * c, const char * what)
{
  if (!c->failed)
  {
    pll_errno = PLL_ERROR_NEWICK_SYNTAX;
    snprintf(pll_errmsg, 200, "newick shim: %s near '%.16s'", what, c->p);
  }
  c->failed = 1;
}

static void nwk_blank(nwk_cursor * c)
{
  while (*c->p == ' ' || *c->p == '\t' || *c->p == '\r' || *c->p == '\n')
    c->p++;
}

static int nwk_stop_char(char ch)
{
  return ch == 0 || ch == '(' || ch == ')' || ch == ',' || ch == ':' ||
         ch == ';' || ch == '[' || ch == ']' || isspace((unsigned char)ch);
}

/* returns a malloc'ed label or NULL when none is present */
static char * nwk_label(nwk_cursor * c)
{
  nwk_blank(c);
  if (*c->p == '\'' || *c->p == '"')
  {
    char q = *c->p++;
    const char * b = c->p;
    while (*c->p && *c->p != q) c->p++;
    if (!*c->p) { nwk_fail(c, "unterminated quoted label"); return NULL; }
    size_t n = (size_t)(c->p - b);
    char * s = (char *)malloc(n + 1);
    memcpy(s, b, n);
    s[n] = 0;
    c->p++;
    return s;
  }
  if (nwk_stop_char(*c->p)) return NULL;
  const char * b = c->p;
  while (!nwk_stop_char(*c->p)) c->p++;
  size_t n = (size_t)(c->p - b);
  char * s = (char *)malloc(n + 1);
  memcpy(s, b, n);
  s[n] = 0;
  return s;
}

/* optional ":" number; absent -> 0 (as the grammar action does) */
static double nwk_length(nwk_cursor * c)
{
  nwk_blank(c);
  if (*c->p != ':') return 0;
  c->p++;
  nwk_blank(c);
  char * e = NULL;
  const char * b = c->p;
  while (!nwk_stop_char(*c->p)) c->p++;
  size_t n = (size_t)(c->p - b);
  char tmp[128];
  if (n == 0 || n >= sizeof(tmp)) { nwk_fail(c, "bad branch length"); return 0; }
  memcpy(tmp, b, n);
  tmp[n] = 0;
  double v = strtod(tmp, &e); /* atof in the grammar == strtod(s, NULL) */
  if (e == tmp) nwk_fail(c, "bad branch length");
  return v;
}

static pll_unode_t * nwk_record(void)
{
  return (pll_unode_t *)calloc(1, sizeof(pll_unode_t));
}

static void nwk_free_graph(pll_unode_t * n)
{
  if (!n) return;
  if (n->next)
  {
    nwk_free_graph(n->next->back);
    nwk_free_graph(n->next->next->back);
    free(n->next->next);
    free(n->next);
  }
  free(n->label);
  free(n);
}

static pll_unode_t * nwk_subtree(nwk_cursor * c)
{
  nwk_blank(c);
  if (*c->p != '(')
  {
    char * lab = nwk_label(c);
    if (!lab) { nwk_fail(c, "expected tip label"); return NULL; }
    pll_unode_t * tip = nwk_record();
    tip->label = lab;
    tip->length = nwk_length(c);
    c->tips++;
    return tip;
  }
  c->p++;
  pll_unode_t * a = nwk_subtree(c);
  nwk_blank(c);
  if (c->failed || *c->p != ',') { nwk_fail(c, "expected ','"); nwk_free_graph(a); return NULL; }
  c->p++;
  pll_unode_t * b = nwk_subtree(c);
  nwk_blank(c);
  if (c->failed || *c->p != ')')
  {
    nwk_fail(c, "inner nodes must be binary");
    nwk_free_graph(a);
    nwk_free_graph(b);
    return NULL;
  }
  c->p++;
  pll_unode_t * up = nwk_record();
  pll_unode_t * l = nwk_record();
  pll_unode_t * r = nwk_record();
  up->next = l; l->next = r; r->next = up;
  l->back = a; a->back = l;
  r->back = b; b->back = r;
  char * lab = nwk_label(c);
  up->label = l->label = r->label = lab;
  up->length = nwk_length(c);
  l->length = a->length;
  r->length = b->length;
  return up;
}

/* parse the whole string into a ring rooted at `top` (already allocated) */
static int nwk_parse(nwk_cursor * c, pll_unode_t * top)
{
  pll_unode_t * s[3] = {NULL, NULL, NULL};
  nwk_blank(c);
  if (*c->p != '(') { nwk_fail(c, "expected '('"); return 0; }
  c->p++;
  for (int i = 0; i < 3 && !c->failed; ++i)
  {
    s[i] = nwk_subtree(c);
    nwk_blank(c);
    char want = (i < 2) ? ',' : ')';
    if (!c->failed && *c->p != want) nwk_fail(c, "top level must be a trifurcation");
    if (!c->failed) c->p++;
  }
  if (c->failed)
  {
    for (int i = 0; i < 3; ++i) nwk_free_graph(s[i]);
    return 0;
  }
  char * lab = nwk_label(c);
  (void)nwk_length(c); /* parsed and dropped, parse_utree.y:202 */
  nwk_blank(c);
  if (c->failed || *c->p != ';')
  {
    nwk_fail(c, "expected ';'");
    free(lab);
    for (int i = 0; i < 3; ++i) nwk_free_graph(s[i]);
    return 0;
  }
  pll_unode_t * b = nwk_record();
  pll_unode_t * d = nwk_record();
  top->next = b; b->next = d; d->next = top;
  top->back = s[0]; s[0]->back = top;
  b->back = s[1];   s[1]->back = b;
  d->back = s[2];   s[2]->back = d;
  top->label = b->label = d->label = lab;
  top->length = s[0]->length;
  b->length = s[1]->length;
  d->length = s[2]->length;
  return 1;
}

/* ---- numbering (parse_utree.y:250-340) --------------------------------- */

typedef struct
{
  unsigned int tip, inner_clv, inner_node;
  int scaler;
} nwk_counters;

static void nwk_number(pll_unode_t * n, nwk_counters * k)
{
  if (!n->next)
  {
    n->node_index = n->clv_index = n->pmatrix_index = k->tip++;
    n->scaler_index = PLL_SCALE_BUFFER_NONE;
    return;
  }
  pll_unode_t * l = n->next;
  pll_unode_t * r = l->next;
  nwk_number(l->back, k);
  nwk_number(r->back, k);
  n->node_index = k->inner_node;
  l->node_index = k->inner_node + 1;
  r->node_index = k->inner_node + 2;
  n->clv_index = l->clv_index = r->clv_index = k->inner_clv;
  n->scaler_index = l->scaler_index = r->scaler_index = k->scaler;
  n->pmatrix_index = k->inner_clv;
  l->pmatrix_index = l->back->pmatrix_index;
  r->pmatrix_index = r->back->pmatrix_index;
  k->inner_clv += 1;
  k->scaler += 1;
  k->inner_node += 3;
}

PLL_EXPORT void pll_utree_reset_template_indices(pll_unode_t * node,
                                                 unsigned int tip_count)
{
  nwk_counters k = {0, tip_count, tip_count, 0};
  pll_unode_t * b = node->next;
  pll_unode_t * d = b->next;
  nwk_number(node->back, &k);
  nwk_number(b->back, &k);
  nwk_number(d->back, &k);
  node->node_index = k.inner_node;
  b->node_index = k.inner_node + 1;
  d->node_index = k.inner_node + 2;
  node->clv_index = b->clv_index = d->clv_index = k.inner_clv;
  node->scaler_index = b->scaler_index = d->scaler_index = k.scaler;
  node->pmatrix_index = node->back->pmatrix_index;
  b->pmatrix_index = b->back->pmatrix_index;
  d->pmatrix_index = d->back->pmatrix_index;
}

/* ---- node table (parse_utree.y:342-445) -------------------------------- */

static void nwk_collect(pll_unode_t * n, pll_unode_t ** tab,
                        unsigned int * tip_at, unsigned int * inner_at)
{
  if (!n->next) { tab[(*tip_at)++] = n; return; }
  nwk_collect(n->next->back, tab, tip_at, inner_at);
  nwk_collect(n->next->next->back, tab, tip_at, inner_at);
  tab[(*inner_at)++] = n;
}

static unsigned int nwk_count(pll_unode_t * n)
{
  if (!n->next) return 1;
  return nwk_count(n->next->back) + nwk_count(n->next->next->back);
}

PLL_EXPORT pll_utree_t * pll_utree_wraptree(pll_unode_t * root,
                                            unsigned int tip_count)
{
  if (tip_count == 0)
  {
    if (!root->next) root = root->back;
    if (!root || !root->next)
    {
      pll_errno = PLL_ERROR_PARAM_INVALID;
      snprintf(pll_errmsg, 200, "Input tree contains no inner nodes.");
      return NULL;
    }
    tip_count = nwk_count(root->back) + nwk_count(root->next->back) +
                nwk_count(root->next->next->back);
  }
  if (tip_count < 3)
  {
    pll_errno = PLL_ERROR_PARAM_INVALID;
    snprintf(pll_errmsg, 200, "Invalid tip_count value (%u).", tip_count);
    return NULL;
  }
  pll_utree_t * t = (pll_utree_t *)malloc(sizeof(pll_utree_t));
  t->nodes = (pll_unode_t **)malloc((2 * tip_count - 2) * sizeof(pll_unode_t *));
  unsigned int tip_at = 0, inner_at = tip_count;
  nwk_collect(root->back, t->nodes, &tip_at, &inner_at);
  nwk_collect(root->next->back, t->nodes, &tip_at, &inner_at);
  nwk_collect(root->next->next->back, t->nodes, &tip_at, &inner_at);
  t->nodes[inner_at] = root;
  t->tip_count = tip_count;
  t->edge_count = 2 * tip_count - 3;
  t->inner_count = tip_count - 2;
  return t;
}

/* ---- teardown (parse_utree.y:37-138) ----------------------------------- */

static void nwk_drop_data(pll_unode_t * n, void (*cb)(void *))
{
  if (n->data && cb) cb(n->data);
}

static void nwk_drop_graph(pll_unode_t * n, void (*cb)(void *))
{
  if (!n) return;
  if (n->next)
  {
    nwk_drop_graph(n->next->back, cb);
    nwk_drop_graph(n->next->next->back, cb);
    nwk_drop_data(n->next, cb);
    nwk_drop_data(n->next->next, cb);
    free(n->next->next);
    free(n->next);
  }
  nwk_drop_data(n, cb);
  free(n->label);
  free(n);
}

PLL_EXPORT void pll_utree_graph_destroy(pll_unode_t * root,
                                        void (*cb_destroy)(void *))
{
  if (!root) return;
  if (root->next && root->back) nwk_drop_graph(root->back, cb_destroy);
  nwk_drop_graph(root, cb_destroy);
}

PLL_EXPORT void pll_utree_destroy(pll_utree_t * tree,
                                  void (*cb_destroy)(void *))
{
  unsigned int total = tree->tip_count + tree->inner_count;
  for (unsigned int i = 0; i < total; ++i)
  {
    pll_unode_t * n = tree->nodes[i];
    nwk_drop_data(n, cb_destroy);
    if (n->next)
    {
      nwk_drop_data(n->next, cb_destroy);
      nwk_drop_data(n->next->next, cb_destroy);
      free(n->next->next);
      free(n->next);
    }
    free(n->label);
    free(n);
  }
  free(tree->nodes);
  free(tree);
}

/* ---- entry points (parse_utree.y:447-526) ------------------------------ */

PLL_EXPORT pll_utree_t * pll_utree_parse_newick_string(const char * s)
{
  nwk_cursor c = {s, 0, 0};
  pll_unode_t * top = nwk_record();
  if (!nwk_parse(&c, top))
  {
    free(top);
    return NULL;
  }
  pll_utree_reset_template_indices(top, c.tips);
  return pll_utree_wraptree(top, c.tips);
}

PLL_EXPORT pll_utree_t * pll_utree_parse_newick(const char * filename)
{
  FILE * f = fopen(filename, "r");
  if (!f)
  {
    pll_errno = PLL_ERROR_FILE_OPEN;
    snprintf(pll_errmsg, 200, "Unable to open file (%s)", filename);
    return NULL;
  }
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  char * buf = (char *)malloc((size_t)n + 1);
  size_t got = fread(buf, 1, (size_t)n, f);
  buf[got] = 0;
  fclose(f);
  pll_utree_t * t = pll_utree_parse_newick_string(buf);
  free(buf);
  return t;
}
