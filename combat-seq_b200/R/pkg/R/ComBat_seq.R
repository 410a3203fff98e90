# ComBat_seq() — drop-in for the reference's one public function (/root/reference/ComBat_seq.R:22-204).
# Same signature, defaults, cat() messages and stop() strings; the numeric body (ComBat_seq.R:33-203) runs in
# libcombatseq.so (hand-written sm_100a CUDA) through .Call("cbs_R_combat_seq") -> cbs_combat_seq() (include/combatseq.h).
# as.matrix / as.factor / model.matrix and the confounding checks stay in R, as in the reference (:25-31, :53-94).
# NOT RUN in the build environment: R is not installed there (see INTEGRATION.md); the Python mirror
# combat-seq_b200/api.py + design.py is the tested twin of this file.
# Extras after the reference's arguments (defaults = reference behaviour): mc.mode, seed (gene.subset.n sampling) and
# n.gpus (0 = every visible GPU; the one R process drives them all through cbs_input.n_gpus).
ComBat_seq <- function(counts, batch, group=NULL, covar_mod=NULL, full_mod=TRUE,
                       shrink=FALSE, shrink.disp=FALSE, gene.subset.n=NULL,
                       mc.mode=c("reference", "stable"), seed=0L, n.gpus=0L){
  counts <- as.matrix(counts)
  dims <- dimnames(counts)
  batch <- as.factor(batch)
  if(any(table(batch)<=1)){
    stop("ComBat-seq doesn't support 1 sample per batch yet")
  }
  n_batch <- nlevels(batch)
  cat("Found", n_batch, "batches\n")
  batchmod <- model.matrix(~-1+batch)
  group <- as.factor(group)
  if(full_mod & nlevels(group)>1){
    cat("Using full model in ComBat-seq.\n")
    mod <- model.matrix(~group)
  }else{
    cat("Using null model in ComBat-seq.\n")
    mod <- model.matrix(~1, data=as.data.frame(t(counts)))
  }
  if(!is.null(covar_mod)){
    if(is.data.frame(covar_mod)){
      covar_mod <- do.call(cbind, lapply(seq_len(ncol(covar_mod)), function(i){model.matrix(~covar_mod[,i])}))
    }
    covar_mod <- covar_mod[, !apply(covar_mod, 2, function(x){all(x==1)}), drop=FALSE]
  }
  mod <- cbind(mod, covar_mod)
  design <- cbind(batchmod, mod)
  check <- apply(design, 2, function(x) all(x == 1))
  design <- as.matrix(design[,!check])
  cat("Adjusting for", ncol(design)-ncol(batchmod), 'covariate(s) or covariate level(s)\n')
  if(qr(design)$rank<ncol(design)){
    if(ncol(design)==(n_batch+1)){stop("The covariate is confounded with batch! Remove the covariate and rerun ComBat-Seq")}
    if(ncol(design)>(n_batch+1)){
      if((qr(design[,-c(1:n_batch)])$rank<ncol(design[,-c(1:n_batch)]))){stop('The covariates are confounded! Please remove one or more of the covariates so the design is not confounded')
      }else{stop("At least one covariate is confounded with batch! Please remove confounded covariates and rerun ComBat-Seq")}}
  }
  NAs <- any(is.na(counts))
  if(NAs){cat(c('Found', sum(is.na(counts)), 'Missing Data Values\n'), sep=' ')}
  # (the library stops on NA / negative / non-integer counts with edgeR::DGEList's wording, as the reference does at :43)
  # mod handed to the library: intercept first, then the non-constant columns (the library rebuilds
  # design = [batch indicators | mod[,-1]], which is cbind(batchmod, mod)[, !check])
  mod_c <- cbind(1, mod[, !apply(mod, 2, function(x) all(x == 1)), drop=FALSE])
  cat("Estimating dispersions\n")
  cat("Fitting the GLM model\n")
  if(shrink){
    cat("Apply shrinkage - computing posterior estimates for parameters\n")
    if(!is.null(gene.subset.n)){cat(sprintf("Using %s random genes for Monte Carlo integration\n", gene.subset.n))
    }else{cat("Using all genes for Monte Carlo integration; the function runs very slow for large number of genes\n")}
    if(!shrink.disp){cat("Apply shrinkage to mean only\n")}
  }else{
    cat("Shrinkage off - using GLM estimates for parameters\n")
  }
  cat("Adjusting the data\n")
  storage.mode(counts) <- "double"; storage.mode(mod_c) <- "double"
  adjust_counts <- .Call("cbs_R_combat_seq", counts, as.integer(batch) - 1L, n_batch, mod_c,
                         as.logical(shrink), as.logical(shrink.disp),
                         if(is.null(gene.subset.n)) -1L else as.integer(gene.subset.n),
                         as.integer(match.arg(mc.mode) == "stable"), as.double(seed), as.integer(n.gpus))
  dimnames(adjust_counts) <- dims
  return(adjust_counts)
}
