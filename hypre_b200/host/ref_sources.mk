# ref_sources.mk -- which files of the reference ($(SRC) = $(REF)/src) make up its host library, shared by
# hypre_b200/host/Makefile (the product's host library: AMG setup stays in the reference's C code, north star)
# and oracle/Makefile (the separately built checker).  The reference's own build system is not run; the
# configuration headers are hypre_b200/host/config{,_mp}/HYPRE_config.h (hand written).
REF_DIRS := utilities blas lapack multivector krylov seq_mv seq_block_mv parcsr_mv \
            parcsr_block_mv distributed_matrix matrix_matrix IJ_mv parcsr_ls \
            distributed_ls/Euclid distributed_ls/ParaSails distributed_ls/pilut

# files that are not part of the host library (Fortran interface wrappers, device
# translation units that are empty without a GPU backend are harmless and kept;
# the rest are stray files that the reference's own per-directory FILES lists omit)
REF_EXCLUDE := %/F90_% %/memory_dmalloc.c %/thread_wrappers.c %/magma.c \
           %/multivector/par_csr_matmultivec.c %/multivector/csr_matmultivec.c \
           %/multivector/par_multivector.c %/multivector/seq_multivector.c %/multivector/backup.c \
           %/parcsr_block_mv/csr_block_matop.c %/IJ_mv/IJMatrix_isis.c %/IJ_mv/IJMatrix_petsc.c \
           %/parcsr_ls/par_sv_interp_lsfit.c \
           %/ParaSails/convert.c %/ParaSails/permute.c \
           %/HYPRE_BuildIJMatrixFromDistributedMatrix.c %/parchord_to_parcsr.c

REF_INCS := -I$(SRC) $(foreach d,$(REF_DIRS),-I$(SRC)/$(d))

ifneq ($(wildcard $(SRC)/HYPRE.h),)
HAVE_REF := 1
REF_ALL  := $(filter-out $(REF_EXCLUDE),$(foreach d,$(REF_DIRS),$(wildcard $(SRC)/$(d)/*.c)))
# stand-alone test mains (driver.c, driver_*.c) are not library code
REF_C    := $(foreach f,$(REF_ALL),$(if $(filter driver%,$(notdir $(f))),,$(f)))
else
HAVE_REF :=
REF_C    :=
endif
