/*
 * orc_octree.c — ORACLE (test infrastructure): ORBextractor::DistributeOctTree
 * (src/ORBextractor.cc:688-1034) and ExtractorNode::DivideNode (:602-674), restated with an
 * explicit doubly-linked list so that list order (push_front / erase) is literal.
 *
 * Canonicalisation: the reference sorts pair<int, ExtractorNode*> (:927), i.e. ties in node size are
 * ordered by heap address.  Here the second key is the node's creation sequence number
 * (ORC_TIE_CANONICAL: ascending, like a bump allocator; ORC_TIE_REVERSED: descending, used only to
 * measure how sensitive the reference's own output is to its allocator).
 */
#include "orc.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct onode {
    int x0, y0, x1, y1;      /* UL=(x0,y0) UR=(x1,y0) BL=(x0,y1) BR=(x1,y1) */
    int* keys; int nkeys;    /* indices into the candidate array, original order preserved */
    int no_more;             /* bNoMore */
    long seq;                /* creation sequence (canonical tie-break) */
    struct onode *prev, *next;
} onode;

typedef struct { onode *head, *tail; int size; long seq; } olist;

static onode* node_new(olist* L, int cap)
{
    onode* n = (onode*)calloc(1, sizeof(onode));
    n->keys = (int*)malloc(sizeof(int) * (size_t)(cap > 0 ? cap : 1));
    n->seq = L->seq++;
    return n;
}
static void node_free(onode* n) { free(n->keys); free(n); }
static void list_push_back(olist* L, onode* n)
{
    n->next = NULL; n->prev = L->tail;
    if (L->tail) L->tail->next = n; else L->head = n;
    L->tail = n; L->size++;
}
static void list_push_front(olist* L, onode* n)
{
    n->prev = NULL; n->next = L->head;
    if (L->head) L->head->prev = n; else L->tail = n;
    L->head = n; L->size++;
}
static onode* list_erase(olist* L, onode* n) /* returns the following element */
{
    onode* nx = n->next;
    if (n->prev) n->prev->next = n->next; else L->head = n->next;
    if (n->next) n->next->prev = n->prev; else L->tail = n->prev;
    L->size--;
    node_free(n);
    return nx;
}

/* DivideNode, :602-674.  Children are created in the order n1..n4 (as ExtractorNode locals they have
 * no list identity yet; sequence numbers are given when pushed, see push_children). */
static void divide_node(const onode* p, const orc_cand* K, onode* c[4], olist* L)
{
    const int halfX = (int)ceilf((float)(p->x1 - p->x0) / 2);   /* :608 */
    const int halfY = (int)ceilf((float)(p->y1 - p->y0) / 2);   /* :609 */
    const int mx = p->x0 + halfX, my = p->y0 + halfY;
    for (int i = 0; i < 4; i++) { c[i] = (onode*)calloc(1, sizeof(onode)); c[i]->keys = (int*)malloc(sizeof(int) * (size_t)p->nkeys); }
    c[0]->x0 = p->x0; c[0]->y0 = p->y0; c[0]->x1 = mx;    c[0]->y1 = my;     /* n1 :613-617 */
    c[1]->x0 = mx;    c[1]->y0 = p->y0; c[1]->x1 = p->x1; c[1]->y1 = my;     /* n2 :621-625 */
    c[2]->x0 = p->x0; c[2]->y0 = my;    c[2]->x1 = mx;    c[2]->y1 = p->y1;  /* n3 :628-632 */
    c[3]->x0 = mx;    c[3]->y0 = my;    c[3]->x1 = p->x1; c[3]->y1 = p->y1;  /* n4 :635-639 */
    for (int i = 0; i < p->nkeys; i++) {                                     /* :643-661 */
        const orc_cand* kp = &K[p->keys[i]];
        int q;
        if (kp->x < mx) q = (kp->y < my) ? 0 : 2;
        else            q = (kp->y < my) ? 1 : 3;
        c[q]->keys[c[q]->nkeys++] = p->keys[i];
    }
    for (int i = 0; i < 4; i++) if (c[i]->nkeys == 1) c[i]->no_more = 1;     /* :665-672 */
    (void)L;
}

typedef struct { int size; long seq; onode* node; } sizeptr;

static int cmp_canon(const void* a, const void* b)
{
    const sizeptr *p = (const sizeptr*)a, *q = (const sizeptr*)b;
    if (p->size != q->size) return p->size < q->size ? -1 : 1;
    return p->seq < q->seq ? -1 : (p->seq > q->seq ? 1 : 0);
}
static int cmp_rev(const void* a, const void* b)
{
    const sizeptr *p = (const sizeptr*)a, *q = (const sizeptr*)b;
    if (p->size != q->size) return p->size < q->size ? -1 : 1;
    return p->seq > q->seq ? -1 : (p->seq < q->seq ? 1 : 0);
}

/* push the non-empty children in order n1..n4 to the FRONT (:826-874 / :935-975); children with more
 * than one key are appended to `vec` (vSizeAndPointerToNode).  Returns number of expandable children. */
static int push_children(olist* L, onode* c[4], sizeptr* vec, int* nvec)
{
    int nexp = 0;
    for (int i = 0; i < 4; i++) {
        if (c[i]->nkeys > 0) {
            c[i]->seq = L->seq++;
            list_push_front(L, c[i]);
            if (c[i]->nkeys > 1) {
                nexp++;
                vec[*nvec].size = c[i]->nkeys; vec[*nvec].seq = c[i]->seq; vec[*nvec].node = c[i];
                (*nvec)++;
            }
        } else {
            node_free(c[i]);
        }
    }
    return nexp;
}

int orc_distribute_octree(const orc_cand* K, int nkeys, int minX, int maxX, int minY, int maxY,
                          int N, int tie_mode, orc_cand* out, int cap)
{
    olist L = {0};
    const int nIni = (int)roundf((float)(maxX - minX) / (maxY - minY));      /* :695 */
    if (nIni < 1) return -1;                                                 /* reference: division by zero */
    const float hX = (float)(maxX - minX) / nIni;                            /* :698 */
    onode** ini = (onode**)malloc(sizeof(onode*) * (size_t)nIni);
    for (int i = 0; i < nIni; i++) {                                         /* :707-729 */
        onode* n = node_new(&L, nkeys);
        n->x0 = (int)(hX * (float)i); n->x1 = (int)(hX * (float)(i + 1));
        n->y0 = 0; n->y1 = maxY - minY;
        list_push_back(&L, n);
        ini[i] = n;
    }
    for (int i = 0; i < nkeys; i++) {                                        /* :733-741 */
        int r = (int)((float)K[i].x / hX);
        ini[r]->keys[ini[r]->nkeys++] = i;
    }
    free(ini);
    for (onode* it = L.head; it;) {                                          /* :746-763 */
        if (it->nkeys == 1) { it->no_more = 1; it = it->next; }
        else if (it->nkeys == 0) it = list_erase(&L, it);
        else it = it->next;
    }

    int finish = 0;
    size_t vcap = (size_t)nkeys * 4 + 64;
    sizeptr* vec = (sizeptr*)malloc(sizeof(sizeptr) * vcap);
    sizeptr* prevvec = (sizeptr*)malloc(sizeof(sizeptr) * vcap);
    int nvec = 0;
    while (!finish) {                                                        /* :780 */
        int prevSize = L.size;
        int nToExpand = 0;
        nvec = 0;
        for (onode* it = L.head; it;) {                                      /* :800-886 */
            if (it->no_more) { it = it->next; continue; }
            onode* c[4];
            divide_node(it, K, c, &L);
            nToExpand += push_children(&L, c, vec, &nvec);
            it = list_erase(&L, it);
        }
        if (L.size >= N || L.size == prevSize) {                             /* :889-895 */
            finish = 1;
        } else if (L.size + nToExpand * 3 > N) {                             /* :909 */
            while (!finish) {
                prevSize = L.size;
                int nprev = nvec;
                memcpy(prevvec, vec, sizeof(sizeptr) * (size_t)nprev);       /* :917 */
                nvec = 0;
                qsort(prevvec, (size_t)nprev, sizeof(sizeptr), tie_mode == ORC_TIE_REVERSED ? cmp_rev : cmp_canon); /* :927 */
                for (int j = nprev - 1; j >= 0; j--) {                       /* :930 */
                    onode* c[4];
                    divide_node(prevvec[j].node, K, c, &L);
                    push_children(&L, c, vec, &nvec);
                    list_erase(&L, prevvec[j].node);                         /* :978 */
                    if (L.size >= N) break;                                  /* :983 */
                }
                if (L.size >= N || L.size == prevSize) finish = 1;           /* :990 */
            }
        }
    }
    free(vec); free(prevvec);

    int nout = 0;
    for (onode* it = L.head; it; it = it->next) {                            /* :1005-1030 */
        int best = it->keys[0];
        int maxResp = K[best].score;
        for (int k = 1; k < it->nkeys; k++)
            if (K[it->keys[k]].score > maxResp) { best = it->keys[k]; maxResp = K[best].score; }
        if (nout < cap) out[nout] = K[best];
        nout++;
    }
    for (onode* it = L.head; it;) it = list_erase(&L, it);
    return nout;
}
