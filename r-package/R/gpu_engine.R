## Drop-in bodies for DiceOptim's criteria on top of the .Call shim (r-package/src/dk_rcall.c).
## Signatures are the reference's (R/EI.R:80, R/EI.grad.R:97, R/qEI.R:119, R/qEI.grad.R:267, R/AKG.R:93,
## R/kriging.quantile.R:92); only the bodies change.  Untested here: the build image has no R.

## Handle cache.  The key hashes everything the device handle depends on (design, response, z, M, trend and covariance
## parameters, diag(T): dkR_model_key, O(n d) per call), so two km objects on the same design with different y / z /
## trend.coef never share a handle.  At most getOption("DiceOptim.gpu.cache", 4) handles are kept; the least recently
## used one is released (device memory freed at once, not at the next gc()).  dk_release() empties the cache.
## Devices: getOption("DiceOptim.gpu.devices", 0L): an integer vector of CUDA ordinals, or NA for every visible GPU;
## with several devices every criterion call splits its rows over them inside the library.
.dk_cache <- new.env(parent = emptyenv())
.dk_cache$handles <- list()     # key -> external pointer, most recently used LAST

.dk_slots <- function(model) {
  cov <- model@covariance
  kern <- match(cov@name, c("gauss", "exp", "matern3_2", "matern5_2", "powexp")) - 1L
  if (is.na(kern)) stop("covariance kernel not supported by the GPU engine: ", cov@name)
  trend <- if (model@trend.formula == ~1) 0L else if (ncol(model@F) == model@d + 1L) 1L else
    stop("only trend formulas ~1 and ~. are supported by the GPU engine")
  list(T = model@T, z = as.numeric(model@z), M = as.matrix(model@M), X = as.matrix(model@X),
       y = as.numeric(model@y), F = as.matrix(model@F), trend.coef = as.numeric(model@trend.coef),
       range.val = as.numeric(cov@range.val),
       shape.val = if (cov@name == "powexp") as.numeric(cov@shape.val) else numeric(0),
       noise.var = if (model@noise.flag) as.numeric(model@noise.var) else numeric(0),
       n = model@n, d = model@d, p = ncol(model@F), kernel = kern, trend = trend, sd2 = cov@sd2,
       nugget = if (cov@nugget.flag) cov@nugget else 0, nugget.flag = cov@nugget.flag)
}

dk_handle <- function(model, devices = getOption("DiceOptim.gpu.devices", 0L)) {
  slots <- .dk_slots(model)
  key <- paste0(.Call("dkR_model_key", slots), "@", paste(devices, collapse = ","))
  hs <- .dk_cache$handles
  h <- hs[[key]]
  if (is.null(h)) {
    keep <- max(1L, as.integer(getOption("DiceOptim.gpu.cache", 4L)))
    while (length(hs) >= keep) { .Call("dkR_release", hs[[1]]); hs[[1]] <- NULL }
    h <- .Call("dkR_model_create", slots, as.integer(devices))
  } else hs[[key]] <- NULL
  hs[[key]] <- h                 # (re)insert at the end: most recently used
  .dk_cache$handles <- hs
  h
}
dk_release <- function() {
  for (h in .dk_cache$handles) .Call("dkR_release", h)
  .dk_cache$handles <- list()
  invisible(NULL)
}
.type_id <- function(type) c(SK = 0L, UK = 1L)[[type]]
.as_rows <- function(x, d) { if (is.null(dim(x))) x <- matrix(as.numeric(x), nrow = 1)
  if (ncol(x) != d) stop("x does not have the right size"); storage.mode(x) <- "double"; as.matrix(x) }

EI <- function(x, model, plugin = NULL, type = "UK", minimization = TRUE, envir = NULL, proxy = FALSE) {
  if (is.null(plugin)) plugin <- if (minimization) min(model@y) else -max(model@y)
  x <- .as_rows(x, model@d)
  r <- .Call("dkR_ei", dk_handle(model), x, plugin, .type_id(type), minimization, proxy, !is.null(envir))
  if (!is.null(envir)) {
    km <- if (minimization) r[[2]] else -r[[2]]
    xcr <- if (proxy) NULL else (plugin - km) / r[[3]]
    assign("xcr", xcr, envir = envir)
    assign("xcr.prob", if (proxy) NULL else pnorm(xcr), envir = envir)
    assign("xcr.dens", if (proxy) NULL else dnorm(xcr), envir = envir)
    assign("kriging.sd", r[[3]], envir = envir)
  }
  r[[1]]
}
EI.grad <- function(x, model, plugin = NULL, type = "UK", minimization = TRUE, envir = NULL, proxy = FALSE) {
  if (is.null(plugin)) plugin <- if (minimization) min(model@y) else -max(model@y)
  g <- .Call("dkR_ei_grad", dk_handle(model), .as_rows(x, model@d), plugin, .type_id(type), minimization, proxy)
  if (is.null(dim(x))) as.numeric(g) else g
}
kriging.quantile <- function(x, model, beta = 0.1, type = "UK", envir = NULL)
  .Call("dkR_quantile", dk_handle(model), .as_rows(x, model@d), qnorm(beta), .type_id(type), FALSE)[[1]]
kriging.quantile.grad <- function(x, model, beta = 0.1, type = "UK", envir = NULL)
  as.numeric(.Call("dkR_quantile", dk_handle(model), .as_rows(x, model@d), qnorm(beta), .type_id(type), TRUE)[[2]])
AKG <- function(x, model, new.noise.var = 0, type = "UK", envir = NULL)
  .Call("dkR_akg", dk_handle(model), .as_rows(x, model@d), new.noise.var, .type_id(type))
qEI <- function(x, model, plugin = NULL, type = "UK", minimization = TRUE, fastCompute = TRUE, eps = 1e-5, envir = NULL) {
  if (!is.matrix(x)) x <- matrix(x, ncol = model@d)
  if (is.null(plugin)) plugin <- if (minimization) min(model@y) else max(model@y)
  .Call("dkR_qei", dk_handle(model), x, nrow(x), plugin, .type_id(type), minimization, FALSE)
}
qEI.grad <- function(x, model, plugin = NULL, type = "UK", minimization = TRUE, fastCompute = TRUE, eps = 1e-6, envir = NULL) {
  if (!is.matrix(x)) x <- matrix(x, ncol = model@d)
  if (is.null(plugin)) plugin <- if (minimization) min(model@y) else -max(model@y)
  if (nrow(x) == 1) return(matrix(EI.grad(x, model, plugin, type), nrow = 1))
  .Call("dkR_qei", dk_handle(model), x, nrow(x), plugin, .type_id(type), minimization, TRUE)
}
## the batched screen of max_crit (R/max_crit.R:104-110) in one call: list(index (1-based), value)
screen_EI <- function(x, model, plugin = NULL, type = "UK", minimization = TRUE, proxy = FALSE, k = 1L) {
  if (is.null(plugin)) plugin <- if (minimization) min(model@y) else -max(model@y)
  .Call("dkR_screen_ei", dk_handle(model), .as_rows(x, model@d), plugin, .type_id(type), minimization, proxy, as.integer(k))
}
## checkPredict (R/checkPredict.R:18-75), vectorised over the rows of x; fastfun members of `model` are skipped as there
checkPredict <- function(x, model, threshold = 1e-4, distance = "covdist", type = "UK") {
  if (is.null(distance)) distance <- "covdist"
  if (is.null(threshold)) threshold <- 1e-4
  kms <- Filter(function(m) is(m, "km"), model)
  x <- .as_rows(x, kms[[1]]@d)
  if (distance == "none") return(rep(FALSE, nrow(x)))
  id <- match(distance, c("euclidean", "covdist", "covratio")) - 1L
  .Call("dkR_check_predict", lapply(kms, dk_handle), x, .type_id(type), id) < threshold
}
## crit_EFI (R/crit_EFI.R:125-228) for every row of x: one upload of x, EI and the Feas factors multiplied on the device
crit_EFI <- function(x, model.fun, model.constraint, equality = FALSE, critcontrol = NULL, plugin = NULL, type = "UK") {
  if (!is(model.constraint, "list")) model.constraint <- list(model.constraint)
  n.cst <- length(model.constraint)
  if (is.null(plugin)) {
    observations <- sapply(model.constraint, function(m) m@y)
    feasibility <- test_feas_vec(cst = observations, equality = equality, tolConstraints = critcontrol$tolConstraints)
    plugin <- if (any(feasibility)) min(model.fun@y[feasibility]) else Inf
  }
  equalityvec <- if (length(equality) == 1 && equality == FALSE) rep(FALSE, n.cst) else equality
  tolvec <- critcontrol$tolConstraints
  if (is.null(tolvec)) { tolvec <- rep(0, n.cst); tolvec[equalityvec] <- 0.05 }
  x <- .as_rows(x, model.fun@d)
  out <- .Call("dkR_feasible_improvement", dk_handle(model.fun), lapply(model.constraint, dk_handle), x, plugin,
               .type_id(type), 0L, as.logical(equalityvec), as.double(tolvec))
  out[checkPredict(x, c(model.constraint, model.fun), type = type, distance = critcontrol$distance,
                   threshold = critcontrol$threshold)] <- 0
  out
}

## ---- the rest of the noisy-criteria family (R/AKG.grad.R:65, R/AEI.R:105, R/AEI.grad.R:65, R/EQI.R:100, R/EQI.grad.R:66) ----
AKG.grad <- function(x, model, new.noise.var = 0, type = "UK", envir = NULL) {
  g <- .Call("dkR_akg_grad", dk_handle(model), .as_rows(x, model@d), new.noise.var, .type_id(type))[[1]]
  if (is.null(dim(x))) as.numeric(g) else g
}
.dk_predict <- function(model, newdata, type = "UK") {
  r <- .Call("dkR_predict", dk_handle(model), .as_rows(newdata, model@d), .type_id(type))
  list(mean = r[[1]], sd = r[[2]])
}
.default_y_min <- function(model, type) {          # R/AEI.R:113-118: mk at the design point minimising mk + qnorm(0.75) sk
  pred <- .dk_predict(model, model@X, type)
  pred$mean[which.min(pred$mean + qnorm(0.75) * pred$sd)]
}
AEI <- function(x, model, new.noise.var = 0, y.min = NULL, type = "UK", envir = NULL) {
  if (is.null(y.min)) y.min <- .default_y_min(model, type)
  .Call("dkR_aei", dk_handle(model), .as_rows(x, model@d), y.min, new.noise.var, .type_id(type), FALSE)
}
AEI.grad <- function(x, model, new.noise.var = 0, y.min = NULL, type = "UK", envir = NULL) {
  if (is.null(y.min)) y.min <- .default_y_min(model, type)
  g <- .Call("dkR_aei", dk_handle(model), .as_rows(x, model@d), y.min, new.noise.var, .type_id(type), TRUE)
  if (is.null(dim(x))) as.numeric(g) else g
}
.default_q_min <- function(model, beta, type) {    # R/EQI.R:108-112: min over the design of mk + qnorm(beta) sk
  pred <- .dk_predict(model, model@X, type)
  min(pred$mean + qnorm(beta) * pred$sd)
}
EQI <- function(x, model, new.noise.var = 0, beta = 0.9, q.min = NULL, type = "UK", envir = NULL) {
  if (is.null(q.min)) q.min <- .default_q_min(model, beta, type)
  .Call("dkR_eqi", dk_handle(model), .as_rows(x, model@d), q.min, new.noise.var, qnorm(beta), .type_id(type), FALSE)
}
EQI.grad <- function(x, model, new.noise.var = 0, beta = 0.9, q.min = NULL, type = "UK", envir = NULL) {
  if (is.null(q.min)) q.min <- .default_q_min(model, beta, type)
  g <- .Call("dkR_eqi", dk_handle(model), .as_rows(x, model@d), q.min, new.noise.var, qnorm(beta), .type_id(type), TRUE)
  if (is.null(dim(x))) as.numeric(g) else g
}

## ---- max_crit (R/max_crit.R:64-130): the random search of EVERY restart goes to the device as ONE matrix ----
## Same arguments and return value.  n.cores is ignored: the forked workers of the reference (mclapply, :125) cannot carry
## a CUDA context; the restarts' candidates are screened in one call (rows r N + 1 .. (r + 1) N belong to restart r) and
## the devices listed in options(DiceOptim.gpu.devices) split the rows.  The L-BFGS-B polish is the reference's optim()
## call on the GPU-backed EI / EI.grad.
max_crit <- function(model, type = "UK", lower, upper, minimization = TRUE, control = NULL, proxy = FALSE, trcontrol = NULL,
                     n.cores = 1) {
  if (is.null(trcontrol$sigma)) trcontrol$sigma <- (upper - lower) / 2
  if (is.null(trcontrol$x0)) trcontrol$x0 <- model@X[which.min(model@y), ]
  if (is.null(trcontrol$crit)) trcontrol$crit <- "EI"
  if (is.null(trcontrol$inTR)) trcontrol$inTR <- FALSE
  if (trcontrol$inTR) {
    lower <- pmax(trcontrol$x0 - trcontrol$sigma, lower)
    upper <- pmin(trcontrol$x0 + trcontrol$sigma, upper)
    if (trcontrol$crit != "EI") proxy <- TRUE
  }
  d <- ncol(model@X)
  if (is.null(control$restarts)) control$restarts <- min(10, d)
  if (is.null(control$BFGS)) control$BFGS <- TRUE
  N <- if (!is.null(control$n.ei)) control$n.ei else min(2e3, 500 * d)
  plugin <- if (minimization) min(model@y) else -max(model@y)
  R <- control$restarts
  x <- matrix(rep(lower, N * R), byrow = TRUE, ncol = d) +
    matrix(runif(d * N * R), ncol = d) * matrix(rep(upper - lower, N * R), byrow = TRUE, ncol = d)
  EI.vect <- EI(x, model, plugin = NULL, type = type, minimization = minimization, envir = NULL, proxy = proxy)
  oo <- lapply(seq_len(R), function(r) {
    rows <- ((r - 1) * N + 1):(r * N)
    I <- which.max(EI.vect[rows])
    o <- list(val = EI.vect[rows][I], par = x[rows[I], , drop = FALSE])
    if (control$BFGS)
      o <- optim(o$par, fn = EI, gr = EI.grad, plugin = plugin, model = model, type = type, minimization = TRUE,
                 envir = NULL, proxy = proxy, method = "L-BFGS-B", lower = lower, upper = upper,
                 control = list(maxit = 100, fnscale = -1, trace = 0), hessian = FALSE)
    o
  })
  allval <- sapply(oo, function(o) if (is.null(o$val)) o$value else o$val)
  I <- which.max(allval)
  list(par = as.numeric(oo[[I]]$par), value = allval[I])
}

## ---- sampleFromEI (R/sampleFromEI.R:77-141): the 1000 d-point screen on the device, the weighted draw by sample() ----
sampleFromEI <- function(model, minimization = TRUE, n = 1, initdistrib = NULL, lower = rep(0, model@d),
                         upper = rep(1, model@d), T = NULL) {
  if (is.null(model)) { message("Error in samplefromEI. A km object needs to be given"); return(NULL) }
  d <- model@d
  if (length(lower) != length(upper)) { message("Error in samplefromEI : lower and upper must have the same length"); return(NULL) }
  initial.points <- if (is.null(initdistrib)) t(lower + t(randLHS(n = d * 1000, k = d)) * (upper - lower)) else initdistrib
  if (d == 1) initial.points <- matrix(initial.points, ncol = 1)
  if (is.null(T)) T <- if (minimization) min(model@y) else max(model@y)
  ei <- .Call("dkR_sample_weights", dk_handle(model), .as_rows(initial.points, d), T, minimization)
  if (sum(ei != 0) < n) {
    maxEI <- max(ei)
    ei <- if (maxEI > 0) ei / maxEI + 1 / (1000 * d) else rep(1, 1000 * d)
  }
  my.points <- initial.points[sample(1:(1000 * d), n, replace = FALSE, prob = ei), ]
  if (d == 1) my.points <- matrix(my.points, ncol = 1)
  if (n == 1) my.points <- matrix(my.points, ncol = d)
  my.points
}

## ---- round-2 entry points ----
.dk_crit <- function(criterion, type = "UK", minimization = TRUE, proxy = FALSE, plugin = 0, qnorm.beta = 0, new.noise.var = 0)
  list(criterion = match(criterion, c("EI", "quantile", "AKG", "AEI", "EQI", "qEI")) - 1L, type = .type_id(type),
       minimization = minimization, proxy = proxy, plugin = plugin, qnorm.beta = qnorm.beta, new.noise.var = new.noise.var)
.dk_seed <- function() as.integer(c(sample.int(.Machine$integer.max, 1), sample.int(.Machine$integer.max, 1)))

## The random search of max_crit with the N candidates drawn on the device (nothing is built in R nor uploaded): list(index, value, x).
## The draws come from the library's counter-based generator seeded from R's RNG state, not from runif() itself.
screen_generated <- function(model, N, lower, upper, criterion = "EI", k = 1L, ...)
  .Call("dkR_screen_generated", dk_handle(model), as.numeric(N), as.numeric(lower), as.numeric(upper), .dk_seed(),
        .dk_crit(criterion, ...), as.integer(k))

## All starts polished together on the device (projected L-BFGS): x0 has one start per row (q rows per start for qEI)
polish <- function(model, x0, lower, upper, criterion = "EI", q = 1L, maxit = 100L, ...)
  .Call("dkR_polish", dk_handle(model), .dk_crit(criterion, ...), as.integer(q), .as_rows(x0, model@d), as.numeric(lower),
        as.numeric(upper), as.integer(maxit))

## max_crit with control$optimizer = "device": same screen, the polish of all restarts in one dk_polish call
max_crit_device <- function(model, type = "UK", lower, upper, minimization = TRUE, control = NULL, proxy = FALSE) {
  d <- ncol(model@X)
  R <- if (is.null(control$restarts)) min(10, d) else control$restarts
  N <- if (!is.null(control$n.ei)) control$n.ei else min(2e3, 500 * d)
  plugin <- if (minimization) min(model@y) else -max(model@y)
  starts <- t(sapply(seq_len(R), function(r)
    screen_generated(model, N, lower, upper, "EI", 1L, type = type, minimization = minimization, proxy = proxy, plugin = plugin)[[3]]))
  o <- polish(model, matrix(starts, ncol = d), lower, upper, "EI", type = type, minimization = TRUE, proxy = proxy, plugin = plugin)
  I <- which.max(o[[2]])
  list(par = as.numeric(o[[1]][I, ]), value = o[[2]][I])
}

## update of the constant-liar loop (R/max_qEI.R:305-316) on the device handle, in place: call it with the model BEFORE the
## rbind and re-key the cache entry to the grown model afterwards (dk_rekey).
dk_append <- function(model, x.new, y.new, new.noise.var = 0)
  invisible(.Call("dkR_model_append", dk_handle(model), as.numeric(x.new), as.numeric(y.new), as.numeric(new.noise.var)))
dk_rekey <- function(old.model, new.model, devices = getOption("DiceOptim.gpu.devices", 0L)) {
  key <- function(m) paste0(.Call("dkR_model_key", .dk_slots(m)), "@", paste(devices, collapse = ","))
  hs <- .dk_cache$handles
  h <- hs[[key(old.model)]]
  if (!is.null(h)) { hs[[key(old.model)]] <- NULL; hs[[key(new.model)]] <- h; .dk_cache$handles <- hs }
  invisible(NULL)
}
