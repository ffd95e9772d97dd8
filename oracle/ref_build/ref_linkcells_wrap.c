/* Compiles the reference linkCells.c where it lies and exports its static
 * getTuple (linkCells.c:497-568) for the cell-numbering parity tests. */
#include "linkCells.c"
void ref_get_tuple(LinkCell* boxes, int iBox, int* t) { getTuple(boxes, iBox, t, t + 1, t + 2); }
