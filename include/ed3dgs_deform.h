/* Deformation-network entry points of ed3dgs_b200 (included by ed3dgs_b200.h). */
#ifndef ED3DGS_DEFORM_H_
#define ED3DGS_DEFORM_H_
#endif
