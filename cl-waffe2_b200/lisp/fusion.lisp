;;;; fusion.lisp — IR rewrite hook + CUDA-graph replay of a compiled iseq + data-parallel helpers.
;;;;
;;;; `on-finalizing-compiling` is called once per `build` per device with the forward / backward
;;;; InstructionSeq (`source/vm/nodes/defnode.lisp:26-57`, `vm/compile.lisp:237-238`; working
;;;; precedent: `source/backends/aten/tensor.lisp:105-141`), AFTER `optimize-memory-locality!` has planned the pool.  A
;;;; rewrite therefore never moves a read or a write to a place where the plan does not expect it: a fused launch sits at
;;;; the position of the LAST instruction of its chain, the chain's other instructions stay in the list with a closure
;;;; that launches nothing (`suppress!`), and a chain is only taken when no foreign launching instruction lies inside it
;;;; (`chain-closed-p`), so every operand still holds what it held for the original instructions.
;;;;
;;;; Rewrites, in the order the method applies them (tests/lisp_fusion_mirror.py restates each one under the same name and
;;;; runs it on recorded instruction lists; tests/test_lisp_fusion_mirror.py):
;;;;   fuse-unfold-chain   Im2ColNode + Permute-Node + the `->contiguous` Move of an unpadded `unfold` (`nn/im2col.lisp:15-54`)
;;;;                       -> wf2cu_unfold
;;;;   fuse-move+arith     Move(copy, x) ; Arith(copy, y) (`!add` = `!copy` + `A+=B`, `base-impl/arithmetic.lisp:150-152`)
;;;;                       -> wf2cu_ew_binary3; save-for-backward places re-aligned to the new argument list
;;;;   fuse-matmul+bias    MatMulNode ; AddNode of LinearLayer's `(!flexible bias)` window (`nn/regression.lisp:14-18`)
;;;;                       -> wf2cu_gemm_ex2, bias in the epilogue
;;;;   fuse-zero-fill+sum  `A*=scal out 0` ; `A+=B out* x` (`base-impl/reduction.lisp:60-79`) -> wf2cu_reduce_sum
;;;;   fuse-softmax-chain  the `!softmax` graph, both `avoid-overflow` forms (`nn/activation.lisp:153-159`) -> wf2cu_softmax_ex
;;;;   fuse-relu           Where-Operation-Node (x > 0) ; Mul (`!relu`, `nn/activation.lisp:26`) -> one fused program
;;;;   fuse-gemm+relu      the gemm above followed by that `!relu` -> act = 1 in the epilogue
;;;;   fold-grad-adder     MatMulNode ; [`!t` window ;] the {GRAD} Move / Add of `expand-gradient-adder`
;;;;                       (`vm/utils.lisp:109-142`) -> the gemm writes / accumulates into the gradient itself
;;;;   alias-immutable-sv4bw  `apply-inst-sv4bw` copies (`vm/vm.lisp:40-65`) whose source is user-owned storage nothing
;;;;                       writes (parameters, the batch) -> the place adopts the source's storage, no copy
;;;; fuse-softmax-chain, fuse-relu and fuse-gemm+relu apply to forward-only graphs (a training graph saves the chain's
;;;; intermediates for backward, and those places must receive what the reference's nodes would put there); the training
;;;; path gets its softmax from the Softmax-Cross-Entropy-Node implementation in composites.lisp.
;;;; `cuda-layer-norm`, `cuda-softmax` and `cuda-self-attention` are offered as drop-in functions for the GPT-2 example's
;;;; plain-function composites (LayerNorm-Revisit, `!softmax` of a pre-divided tensor, SelfAttention's score block).
(in-package :cl-waffe2/backends.cuda)

;; ---- bookkeeping -----------------------------------------------------------------------------------------------------
(defvar *rewritten* (trivial-garbage:make-weak-hash-table :weakness :key :test #'eq)
  "WfInstruction -> what its closure does now: (:identity) (:binary3 code) (:gemm ta tb bias-p act) (:reduce-sum) (:relu) ...
An instruction without an entry still runs its node's own implementation.")

(defun rewritten (inst) (gethash inst *rewritten*))
(defun (setf rewritten) (what inst) (setf (gethash inst *rewritten*) what))

(defun raw-node-p (inst type)
  "INST still runs the implementation of a node of TYPE."
  (and (null (rewritten inst)) (typep (wf/vm:wfop-node inst) type)))

(defun rewritten-as-p (inst kind &optional (code nil code-p))
  (let ((what (rewritten inst)))
    (and what (eq (car what) kind) (or (not code-p) (eql (second what) code)))))

(defun window-inst-p (inst)
  "Host-side pointer work only: `!view`, `!permute`, `!t`."
  (typep (wf/vm:wfop-node inst) '(or ViewTensorNode cl-waffe2/base-impl::Permute-Node LazyTransposeNode)))

(defun suppressed-p (inst) (rewritten-as-p inst :identity))

(defun no-sv4bw-p (inst) (every #'null (wf/vm:wfop-sv4bw inst)))

(defun cuda-inst-p (inst)
  (every #'(lambda (x) (typep x 'CUDATensor)) (wf/vm:wfop-args inst)))

(defun suppress! (inst)
  "INST keeps its place, `out-to` and bookkeeping but launches nothing: its closure returns the argument `out-to` aliases
(first for arithmetic / Move, third for MatMulNode, the OUT operand of `(x out)` nodes otherwise)."
  (let ((k (typecase (wf/vm:wfop-node inst)
	     ((or AddNode SubNode MulNode DivNode MoveTensorNode) 0)
	     (MatMulNode 2)
	     (t 1))))
    (setf (rewritten inst) '(:identity)
	  (wf/vm:wfop-op inst) #'(lambda (&rest args) (nth k args)))))

(defun next-launching (iseq)
  "(values skipped inst rest): INST is the first instruction of ISEQ that is neither suppressed nor a window (the
topological order puts the window of a second operand between a producer and its consumer), SKIPPED those before it."
  (loop with skipped = nil
	for tail on iseq
	for inst = (car tail)
	if (or (window-inst-p inst) (suppressed-p inst))
	  do (push inst skipped)
	else
	  do (return (values (nreverse skipped) inst (cdr tail)))
	finally (return (values (nreverse skipped) nil nil))))

(defun producer-table (iseq)
  "out-to tensor -> the instruction that produces it."
  (let ((table (make-hash-table :test #'eq)))
    (dolist (inst iseq table)
      (dolist (o (wf/vm:wfop-out-to inst))
	(setf (gethash o table) inst)))))

(defun use-table (iseq)
  "tensor -> how many instructions that still run read it."
  (let ((table (make-hash-table :test #'eq)))
    (dolist (inst iseq table)
      (unless (suppressed-p inst)
	(dolist (a (wf/vm:wfop-args inst))
	  (incf (gethash a table 0)))))))

(defun bcast-axes (tensor)
  (loop for v in (tensor-view tensor)
	for k fixnum upfrom 0
	when (subscript-broadcast v) collect k))

(defun through-views (tensor producers)
  "Follow ViewTensorNodes (args: viewed-tensor old, `base-impl/fundamental.lisp:173-184`) back to the tensor they window."
  (loop for p = (gethash tensor producers)
	while (and p (typep (wf/vm:wfop-node p) 'ViewTensorNode))
	do (setq tensor (second (wf/vm:wfop-args p)))
	finally (return tensor)))

(defun constant-value (tensor producers)
  "The number when TENSOR is (a broadcast window of) a one-element tensor made at build time, else NIL."
  (let ((base (through-views tensor producers)))
    (and (not (cl-waffe2/vm::tensor-tmp-p base))
	 (not (scalar-p base))
	 (every #'(lambda (s) (eql s 1)) (shape base))
	 (ignore-errors (tensor-vec base) (vref base 0)))))

;; ---- fused element-wise programs (include/waffe2cuda.h: WF2CU_P_*; 4 x int32 per instruction) --------------------------
;; csrc/fused.cu compiles the programs the hosts emit (its catalogue lists each one next to the reference lines it restates);
;; anything else runs on the same entry point's interpreter.
(defconstant +p-load+ 0) (defconstant +p-store+ 1) (defconstant +p-imm+ 2) (defconstant +p-add+ 3) (defconstant +p-sub+ 4)
(defconstant +p-mul+ 5)  (defconstant +p-div+ 6)   (defconstant +p-unary+ 7)
(defconstant +p-gt+ (logior 9 (ash 3 8)))	; WF2CU_P_CMPSEL | (WF2CU_GT << 8): r[dst] = r[a] > r[b] ? 1 : 0
(defconstant +u-loge+ 11)			; WF2CU_LOGE

(defun float-bits (x)
  "The IEEE-754 bits of the single-float X as a signed 32-bit integer (the immediate field of WF2CU_P_IMM)."
  (let ((u (ieee-floats:encode-float32 (coerce x 'single-float))))
    (if (>= u #x80000000) (- u #x100000000) u)))

(defun run-fused-program (program tensors)
  "PROGRAM: list of (opcode dst a b); TENSORS: equal-shape CUDATensors, indexed by the LOAD / STORE instructions."
  (let ((n (length program))
	(nviews (length tensors)))
    (with-foreign-objects ((prog :int32 (* 4 n))
			   (views '(:struct wf2cu-view) nviews))
      (loop for ins in program
	    for k fixnum upfrom 0
	    do (loop for field in ins
		     for f fixnum upfrom 0
		     do (setf (mem-aref prog :int32 (+ (* 4 k) f)) field)))
      (loop for tensor in tensors
	    for j fixnum upfrom 0
	    do (tensor-vec tensor)	; make sure storage exists (lazy allocation)
	       (lower-view! tensor (mem-aptr views '(:struct wf2cu-view) j)))
      (fused prog n views nviews))))

;; ---- Move + Arith ----------------------------------------------------------------------------------------------------------
(defun arith-op-code (node)
  (typecase node
    (AddNode +add+) (SubNode +sub+) (MulNode +mul+) (DivNode +div+) (t nil)))

(defun fuse-move+arith (iseq)
  "copy <- MoveTensor(copy, x) ; copy <- Arith(copy, y)   ==>   copy <- binary3(op, copy, x, y).
The Arith's save-for-backward places follow its arguments: (p_copy p_y) -> (nil p_copy p_y) — `copy` held x when
`apply-inst-sv4bw` would have saved it, so saving x into p_copy stores the same values."
  (loop with out = nil
	while iseq
	do (let ((i0 (pop iseq)))
	     (multiple-value-bind (skipped i1 rest) (next-launching iseq)
	       (if (and i1
			(raw-node-p i0 'MoveTensorNode)
			(null (rewritten i1))
			(arith-op-code (wf/vm:wfop-node i1))
			(cuda-inst-p i0) (cuda-inst-p i1)
			(not (movetensor-ignore-me (wf/vm:wfop-node i0)))
			(not (cl-waffe2/vm::wfop-grad-adder-p i0))
			(no-sv4bw-p i0)
			;; the arith node works in place on the Move's destination, and is not a reduction
			(eq (first (wf/vm:wfop-out-to i0)) (first (wf/vm:wfop-args i1)))
			(null (bcast-axes (first (wf/vm:wfop-args i1)))))
		   (let ((code (arith-op-code (wf/vm:wfop-node i1))))
		     ;; keep i1's bookkeeping (out-to, node, self), replace its closure and arguments
		     (setf (wf/vm:wfop-args i1) (list (first (wf/vm:wfop-args i0))
						      (second (wf/vm:wfop-args i0))
						      (second (wf/vm:wfop-args i1)))
			   (wf/vm:wfop-sv4bw i1) (and (wf/vm:wfop-sv4bw i1)
						      (list nil (first (wf/vm:wfop-sv4bw i1)) (second (wf/vm:wfop-sv4bw i1))))
			   (rewritten i1) (list :binary3 code)
			   (wf/vm:wfop-op i1)
			   #'(lambda (dst x y)
			       (with-lowered-views ((dv dst) (xv x) (yv y))
				 (ew-binary3 code dv xv yv))
			       dst))
		     (dolist (s skipped) (push s out))
		     (push i1 out)
		     (setq iseq rest))
		   (push i0 out))))
	finally (return (nreverse out))))

;; ---- MatMul + bias [+ relu] ---------------------------------------------------------------------------------------------------
(defun bias-operand-p (tensor n)
  "T when TENSOR is a rank-2 view whose row axis is (:broadcast _) over a contiguous N-element vector:
what `(!flexible bias)` becomes inside `step-linear` (`nn/regression.lisp:14-18`)."
  (and (typep tensor 'CUDATensor)
       (= (dims tensor) 2)
       (subscript-broadcast (first (tensor-view tensor)))
       (not (subscript-broadcast (second (tensor-view tensor))))
       (= (observe (second (shape tensor))) n)
       (eql (observe (second (tensor-stride tensor))) 1)))

(defun launch-gemm (ta tb a b c &key bias (act 0) residual feeds-matmul)
  "C <- op(A) op(B) [+ bias] [act] [+ residual] on rank-2 views; TA / TB: the operand is read untransposed and flagged, as
MatMulNode's implementation does (`read-untransposed`, matrix-ops.lisp:52-56).
FEEDS-MATMUL: C is the untransposed left operand of the product issued next (the two MLP projections of
`examples/gpt-2/model.lisp:93-137`): the epilogue also writes that product's tensor-core operand plane
(wf2cu_emit_operand_planes, one shot; honoured by the 256-row kernels, ignored otherwise; same bits as the pre-pass)."
  (when feeds-matmul (emit-operand-planes 1))
  (let* ((a* (read-untransposed a))
	 (b* (read-untransposed b))
	 (m (observe (first (shape c)))) (n (observe (second (shape c))))
	 (k (observe (if ta (first (shape a*)) (second (shape a*))))))
    (with-lowered-views ((av a*) (bv b*) (cv c))
      (flet ((off (v) (foreign-slot-value v '(:struct wf2cu-view) 'offset))
	     (ld  (v) (mem-aref (foreign-slot-pointer v '(:struct wf2cu-view) 'stride) :int64 0))
	     (ptr (x) (device-vec-ptr (tensor-vec x))))
	(if bias
	    (with-lowered-views ((biasv bias))
	      (if residual
		  (with-lowered-views ((rv residual))
		    (gemm-ex2 (if ta 1 0) (if tb 1 0) m n k (ptr a*) (off av) (ld av) (ptr b*) (off bv) (ld bv) (ptr c) (off cv) (ld cv)
			      (ptr bias) (off biasv) act (ptr residual) (off rv) (ld rv)))
		  (gemm-ex2 (if ta 1 0) (if tb 1 0) m n k (ptr a*) (off av) (ld av) (ptr b*) (off bv) (ld bv) (ptr c) (off cv) (ld cv)
			    (ptr bias) (off biasv) act (null-pointer) 0 0)))
	    (if residual
		(with-lowered-views ((rv residual))
		  (gemm-ex2 (if ta 1 0) (if tb 1 0) m n k (ptr a*) (off av) (ld av) (ptr b*) (off bv) (ld bv) (ptr c) (off cv) (ld cv)
			    (null-pointer) 0 act (ptr residual) (off rv) (ld rv)))
		(gemm-ex2 (if ta 1 0) (if tb 1 0) m n k (ptr a*) (off av) (ld av) (ptr b*) (off bv) (ld bv) (ptr c) (off cv) (ld cv)
			  (null-pointer) 0 act (null-pointer) 0 0)))))
    c))

(defun fuse-matmul+bias (iseq)
  "out <- MatMul(a, b, out) ; out <- Add(out, bias*)   ==>   out <- gemm_ex2(a, b, out, bias).  The Add instruction is
kept (its `out-to` is what the rest of the graph reads) with the MatMul's arguments and save-for-backward places."
  (loop with result = nil
	while iseq
	do (let ((i0 (pop iseq)))
	     (multiple-value-bind (skipped i1 rest) (next-launching iseq)
	       (if (and i1
			(raw-node-p i0 'MatMulNode)
			(raw-node-p i1 'AddNode)
			(cuda-inst-p i0) (cuda-inst-p i1)
			(no-sv4bw-p i1)
			(= (dims (third (wf/vm:wfop-args i0))) 2)
			(eq (first (wf/vm:wfop-out-to i0)) (first (wf/vm:wfop-args i1)))
			(bias-operand-p (second (wf/vm:wfop-args i1))
					(observe (second (shape (third (wf/vm:wfop-args i0)))))))
		   (let* ((node (wf/vm:wfop-node i0))
			  (ta (trans-a? node)) (tb (trans-b? node))
			  (bias (second (wf/vm:wfop-args i1))))
		     (setf (wf/vm:wfop-args i1) (append (wf/vm:wfop-args i0) (list bias))
			   (wf/vm:wfop-sv4bw i1) (and (wf/vm:wfop-sv4bw i0) (append (wf/vm:wfop-sv4bw i0) (list nil)))
			   (rewritten i1) (list :gemm ta tb t 0)
			   (wf/vm:wfop-op i1)
			   #'(lambda (a b out bias)
			       (launch-gemm ta tb a b out :bias bias)))
		     (dolist (s skipped) (push s result))
		     (push i1 result)
		     (setq iseq rest))
		   (push i0 result))))
	finally (return (nreverse result))))

(defun fuse-gemm+relu (iseq)
  "out <- gemm(a, b, out, bias) ; dst <- relu(dst, out)   ==>   dst <- gemm with act = 1, when nothing else reads OUT."
  (loop with uses = (use-table iseq)
	with result = nil
	while iseq
	do (let ((i0 (pop iseq)))
	     (multiple-value-bind (skipped i1 rest) (next-launching iseq)
	       (if (and i1
			(rewritten-as-p i0 :gemm)
			(eql (fifth (rewritten i0)) 0)
			(rewritten-as-p i1 :relu)
			(eq (second (wf/vm:wfop-args i1)) (first (wf/vm:wfop-out-to i0)))
			(eql (gethash (first (wf/vm:wfop-out-to i0)) uses) 1)
			(no-sv4bw-p i0) (no-sv4bw-p i1))
		   (destructuring-bind (ta tb bias-p act) (cdr (rewritten i0))
		     (declare (ignore act))
		     (let ((dst (first (wf/vm:wfop-args i1))))
		       ;; arguments: the gemm's, then the tensor the relu wrote
		       (setf (wf/vm:wfop-args i1) (append (wf/vm:wfop-args i0) (list dst))
			     (wf/vm:wfop-sv4bw i1) nil
			     (rewritten i1) (list :gemm ta tb bias-p 1)
			     (wf/vm:wfop-op i1)
			     (if bias-p
				 #'(lambda (a b out bias dst) (declare (ignore out)) (launch-gemm ta tb a b dst :bias bias :act 1))
				 #'(lambda (a b out dst) (declare (ignore out)) (launch-gemm ta tb a b dst :act 1)))))
		     (dolist (s skipped) (push s result))
		     (push i1 result)
		     (setq iseq rest))
		   (push i0 result))))
	finally (return (nreverse result))))

;; ---- `!sum`: zero fill + A+=B into a broadcast window -------------------------------------------------------------------------
(defun fuse-zero-fill+sum (iseq)
  "out <- Mul(out, 0*) ; out* <- View(out*, out) ; out* <- Add(out*, x)   ==>   out* <- reduce_sum(x), no zero fill
(`!sum`, base-impl/reduction.lisp:60-79: `A*=scal out 0` exists only to clear the accumulator).
Returns (values iseq sums); SUMS maps the Add's out-to tensor to (x add-instruction zero-fill-instruction)."
  (let ((producers (producer-table iseq))
	(uses (use-table iseq))
	(sums (make-hash-table :test #'eq)))
    (dolist (ia iseq)
      (when (and (raw-node-p ia 'AddNode) (cuda-inst-p ia) (no-sv4bw-p ia)
		 (bcast-axes (first (wf/vm:wfop-args ia))))
	(let* ((window (first (wf/vm:wfop-args ia)))
	       (pv (gethash window producers)))
	  (when (and pv (typep (wf/vm:wfop-node pv) 'ViewTensorNode) (eql (gethash window uses) 1))
	    (let* ((acc (second (wf/vm:wfop-args pv)))
		   (pz  (gethash acc producers)))
	      (when (and pz (raw-node-p pz 'MulNode) (no-sv4bw-p pz)
			 (eql (gethash acc uses) 1)
			 (eql (constant-value (second (wf/vm:wfop-args pz)) producers) 0.0))
		(suppress! pz)
		(setf (rewritten ia) '(:reduce-sum)
		      (wf/vm:wfop-op ia)
		      #'(lambda (out* x)
			  (with-lowered-views ((xv x) (ov out*))
			    (reduce-sum xv ov 0 1.0))
			  out*)
		      (gethash (first (wf/vm:wfop-out-to ia)) sums)
		      (list (second (wf/vm:wfop-args ia)) ia pz))))))))
    (values iseq sums)))

(defun sum-source (tensor producers sums)
  "The (x add-instruction zero-fill-instruction) entry when TENSOR is, through windows only, the result of a reduce-sum."
  (loop
    (let ((entry (gethash tensor sums)))
      (when entry (return entry)))
    (let ((p (gethash tensor producers)))
      (if (and p (typep (wf/vm:wfop-node p) 'ViewTensorNode))
	  (setq tensor (second (wf/vm:wfop-args p)))
	  (return nil)))))

(defun chain-closed-p (iseq members result)
  "MEMBERS occupy consecutive positions once windows are ignored, none of them saves anything for backward, and nothing
they produce (or a window of it) is read by an instruction outside the chain, except RESULT."
  (let* ((positions (sort (mapcar #'(lambda (m) (position m iseq :test #'eq)) members) #'<))
	 (made (make-hash-table :test #'eq)))
    (and (every #'no-sv4bw-p members)
	 (loop for p from (first positions) to (car (last positions))
	       for inst = (nth p iseq)
	       always (or (member inst members :test #'eq) (window-inst-p inst)))
	 (progn
	   (dolist (m members)
	     (dolist (o (wf/vm:wfop-out-to m))
	       (unless (eq o result) (setf (gethash o made) t))))
	   (dolist (inst iseq)
	     (when (and (window-inst-p inst)
			(some #'(lambda (a) (gethash a made)) (wf/vm:wfop-args inst)))
	       (dolist (o (wf/vm:wfop-out-to inst)) (setf (gethash o made) t))))
	   (loop for inst in iseq
		 never (and (not (member inst members :test #'eq))
			    (not (window-inst-p inst))
			    (some #'(lambda (a) (gethash a made)) (wf/vm:wfop-args inst))))))))

;; ---- `!softmax` -----------------------------------------------------------------------------------------------------------------
(defun fuse-softmax-chain (iseq sums)
  "The `!softmax` graph (nn/activation.lisp:153-159) after fuse-move+arith and fuse-zero-fill+sum:
     e <- Exp(x1, copy) ; s <- reduce_sum(e) ; out <- binary3(Div, copy2, e, s)
   with x1 = x (avoid-overflow nil), or x1 <- binary3(Sub, copy0, x, mean*) where mean* windows
   binary3(Div, copy1, reduce_sum(x), n*) over the same axis and n = the axis' length (avoid-overflow t: `!mean`,
   base-impl/reduction.lisp:81-136)
   ==> ONE wf2cu_softmax_ex at the position of the final Div (it reads x and writes the Div's destination; when the pool
   gave both the same storage the kernel runs in place: every element is read and written by the same thread, after the
   row statistics are complete); everything before it in the chain is suppressed."
  (let ((producers (producer-table iseq)))
    (dolist (idiv iseq)
      (when (rewritten-as-p idiv :binary3 +div+)
	(destructuring-bind (dst e s) (wf/vm:wfop-args idiv)
	  (let* ((entry (gethash s sums))
		 (iexp  (gethash e producers)))
	    (when (and entry iexp
		       (eq (first entry) e)
		       (raw-node-p iexp 'ExpNode)
		       (= (length (bcast-axes s)) 1))
	      (let* ((axis (first (bcast-axes s)))
		     (members (list iexp (second entry) (third entry) idiv))
		     (icopy (gethash (second (wf/vm:wfop-args iexp)) producers))
		     (x1 (first (wf/vm:wfop-args iexp)))
		     (avoid-overflow nil))
		;; the (!copy x) ExpNode writes into
		(when (and icopy (raw-node-p icopy 'MoveTensorNode))
		  (push icopy members))
		(let ((isub (gethash x1 producers)))
		  (when (and isub (rewritten-as-p isub :binary3 +sub+))
		    (destructuring-bind (copy0 x0 mean*) (wf/vm:wfop-args isub)
		      (declare (ignore copy0))
		      (let ((imean (gethash (through-views mean* producers) producers)))
			(when (and imean (rewritten-as-p imean :binary3 +div+)
				   (equal (bcast-axes mean*) (list axis)))
			  (let ((msum (sum-source (second (wf/vm:wfop-args imean)) producers sums))
				(n    (constant-value (third (wf/vm:wfop-args imean)) producers)))
			    (when (and msum n
				       (eq (first msum) x0)
				       (equal (bcast-axes (first (wf/vm:wfop-args (second msum)))) (list axis))
				       (= n (observe (nth axis (shape x0)))))
			      (setq members (append members (list isub imean (second msum) (third msum)))
				    x1 x0
				    avoid-overflow t))))))))
		(when (chain-closed-p iseq members (first (wf/vm:wfop-out-to idiv)))
		  (dolist (m members)
		    (unless (eq m idiv) (suppress! m)))
		  (let ((avoid (if avoid-overflow 1 0)))
		    (setf (wf/vm:wfop-args idiv) (list dst x1)
			  (rewritten idiv) (list :softmax axis avoid-overflow)
			  (wf/vm:wfop-op idiv)
			  #'(lambda (dst x)
			      (with-lowered-views ((xv x) (dv dst))
				(softmax-ex xv dv axis avoid 0.0))
			      dst))))))))))
    iseq))

;; ---- `!relu` ------------------------------------------------------------------------------------------------------------------------
(defun fuse-relu (iseq)
  "mask <- Where(x > 0 ? 1 : 0) ; out <- binary3(Mul, copy, x, mask)   ==>   out <- x * (x > 0 ? 1 : 0) as one fused program,
when nothing else reads the mask and nothing is saved for backward (`!relu`, nn/activation.lisp:26)."
  (let ((producers (producer-table iseq)))
    (dolist (imul iseq)
      (when (rewritten-as-p imul :binary3 +mul+)
	(destructuring-bind (dst x mask) (wf/vm:wfop-args imul)
	  (let ((iw (gethash mask producers)))
	    (when (and iw (raw-node-p iw 'Where-Operation-Node)
		       (eq (first (wf/vm:wfop-args iw)) x)
		       ;; `A>scal x 0.0` records (list #'> 0.0) (base-impl/logical.lisp:185-189)
		       (let ((info (logical-compiler-info (wf/vm:wfop-node iw))))
			 (and info (eql (car info) #'>) (realp (second info)) (zerop (second info))))
		       (= (logical-true-then (wf/vm:wfop-node iw)) 1)
		       (= (logical-false-then (wf/vm:wfop-node iw)) 0)
		       (chain-closed-p iseq (list iw imul) (first (wf/vm:wfop-out-to imul))))
	      (suppress! iw)
	      (setf (wf/vm:wfop-args imul) (list dst x)
		    (rewritten imul) '(:relu)
		    (wf/vm:wfop-op imul)
		    #'(lambda (dst x)
			;; csrc/fused.cu PRelu
			(run-fused-program
			 `((,+p-load+ 0 0 0) (,+p-imm+ 1 0 ,(float-bits 0.0)) (,+p-gt+ 2 0 1) (,+p-mul+ 2 0 2) (,+p-store+ 2 1 0))
			 (list x dst))
			dst)))))))
    iseq))

;; ---- gradient adders (`expand-gradient-adder`, vm/utils.lisp:109-142) -------------------------------------------------------------
(defun fold-grad-adder (iseq)
  "g <- MatMul(a, b, g) ; [gp <- Permute(g) ;] grad <- Move|Add(grad, g|gp) with grad-adder-p
   ==>  one gemm that writes (Move: first gradient) or accumulates into (Add) the gradient itself.  Through the window `!t`
   leaves behind (LinearLayer's weight is used as (!t weight), so its gradient arrives transposed) the product is formed
   transposed: (op(a) op(b))^T = op(b)^T op(a)^T — operands swapped, both flags inverted.  Gradients are user-owned storage
   (not pool memory) and the instructions are adjacent, so writing the gradient one instruction early is not observable."
  (loop for tail on iseq
	for i0 = (car tail)
	when (and (raw-node-p i0 'MatMulNode) (cuda-inst-p i0) (no-sv4bw-p i0)
		  (= (dims (third (wf/vm:wfop-args i0))) 2))
	  do (let* ((i1 (second tail))
		    (perm (and i1 (raw-node-p i1 'cl-waffe2/base-impl::Permute-Node)
			       (eq (first (wf/vm:wfop-args i1)) (first (wf/vm:wfop-out-to i0)))
			       i1))
		    (adder (if perm (third tail) i1))
		    (src (first (wf/vm:wfop-out-to (or perm i0)))))
	       (when (and adder
			  (cl-waffe2/vm::wfop-grad-adder-p adder)
			  (null (rewritten adder))
			  (typep (wf/vm:wfop-node adder) '(or MoveTensorNode AddNode))
			  (eq (second (wf/vm:wfop-args adder)) src)
			  (typep (first (wf/vm:wfop-args adder)) 'CUDATensor)
			  (null (bcast-axes (first (wf/vm:wfop-args adder))))
			  ;; a `!t` window is exactly the last two axes exchanged
			  (or (null perm)
			      (equal (shape (first (wf/vm:wfop-args adder)))
				     (reverse (shape (third (wf/vm:wfop-args i0)))))))
		 (let* ((node (wf/vm:wfop-node i0))
			(ta (trans-a? node)) (tb (trans-b? node))
			(accumulate (typep (wf/vm:wfop-node adder) 'AddNode))
			(swap (and perm t))
			(grad (first (wf/vm:wfop-args adder))))
		   (setf (rewritten i0) (list :gemm (if swap (not tb) ta) (if swap (not ta) tb) nil 0)
			 (wf/vm:wfop-op i0)
			 #'(lambda (a b out)
			     (if swap
				 (launch-gemm (not tb) (not ta) b a grad :residual (and accumulate grad))
				 (launch-gemm ta tb a b grad :residual (and accumulate grad)))
			     out))
		   (suppress! adder)))))
  iseq)

;; ---- save-for-backward copies -----------------------------------------------------------------------------------------------------
(defun raw-vec (tensor)
  "TENSOR's storage without forcing a lazy allocation."
  (slot-value tensor 'cl-waffe2/vm.generic-tensor::vec))

(defun alias-immutable-sv4bw (iseq-fw iseq-bw)
  "`apply-inst-sv4bw` (vm/vm.lisp:40-65) copies every saved argument into its place before the instruction runs, every step.
When the argument is user-owned storage (not pool memory: a parameter, a constant, the batch) that no instruction of either
sequence writes, the copy is dropped and the place adopts the argument's storage as the instruction runs — what the
reference itself does for scalar places (`(setf (tensor-vec place) (tensor-vec var))`).  Runs last: the places are aligned
with the final argument lists."
  (let ((written nil))
    (dolist (inst (append iseq-fw iseq-bw))
      (unless (window-inst-p inst)
	(dolist (o (wf/vm:wfop-out-to inst))
	  (let ((v (raw-vec o)))
	    (when v (pushnew v written :test #'eq))))))
    (dolist (inst iseq-fw)
      (let ((adopt nil))
	(loop for place in (wf/vm:wfop-sv4bw inst)
	      for var in (wf/vm:wfop-args inst)
	      for k fixnum upfrom 0
	      when (and place var
			(typep var 'CUDATensor)
			(not (scalar-p var)) (not (scalar-p place))
			(not (cl-waffe2/vm::tensor-tmp-p var))
			(raw-vec var)
			(not (member (raw-vec var) written :test #'eq)))
		do (push (cons k place) adopt))
	(when adopt
	  (setf (wf/vm:wfop-sv4bw inst)
		(loop for place in (wf/vm:wfop-sv4bw inst)
		      for k fixnum upfrom 0
		      collect (if (assoc k adopt) nil place)))
	  (let ((op (wf/vm:wfop-op inst)))
	    (setf (wf/vm:wfop-op inst)
		  #'(lambda (&rest args)
		      (loop for (k . place) in adopt
			    do (setf (tensor-vec place) (tensor-vec (nth k args))))
		      (apply op args)))))))
    (values iseq-fw iseq-bw)))

;; ---- unfold: Im2ColNode ; Permute-Node ; MoveTensor(->contiguous)  ==>  one wf2cu_unfold
;; `!im2col` (`nn/im2col.lisp:15-54`) always emits  col <- Im2Col(x, col) ; p <- Permute(col) ;
;; dense <- MoveTensor(dense, p)  (the `->contiguous` inside `!reshape`).  With padding (0 0) — MNIST-CNN,
;; the pooling layers — the three instructions become one launch that reads the image through its view
;; and writes the (N*Ho*Wo, C*kh*kw) rows directly.  With padding the chain is left as it is (the padded
;; image is produced by a Move into a window, and the node then skips trailing rows: three launches).
(defun unpadded-unfold-p (node)
  (with-slots (padding-h padding-w dilation-h dilation-w) node
    (and (eql padding-h 0) (eql padding-w 0) (eql dilation-h 1) (eql dilation-w 1))))

(defun fuse-unfold-chain (iseq)
  (loop with out = nil
	while iseq
	do (let* ((i0 (pop iseq))
		  (i1 (first iseq))
		  (i2 (second iseq)))
	     (if (and i1 i2
		      (typep (wf/vm:wfop-node i0) 'Im2ColNode)
		      (unpadded-unfold-p (wf/vm:wfop-node i0))
		      (typep (wf/vm:wfop-node i1) 'cl-waffe2/base-impl::Permute-Node)
		      (typep (wf/vm:wfop-node i2) 'MoveTensorNode)
		      (cuda-inst-p i0) (cuda-inst-p i2)
		      (null (wf/vm:wfop-sv4bw i0))
		      (eq (first (wf/vm:wfop-out-to i0)) (first (wf/vm:wfop-args i1)))
		      (eq (first (wf/vm:wfop-out-to i1)) (second (wf/vm:wfop-args i2))))
		 (let ((node (wf/vm:wfop-node i0))
		       (x    (first (wf/vm:wfop-args i0))))
		   (pop iseq) (pop iseq)
		   (setf (wf/vm:wfop-args i2) (list (first (wf/vm:wfop-args i2)) x)
			 (wf/vm:wfop-op i2)
			 #'(lambda (dense image)
			     (with-unfold-geom (g node (shape image))
			       (with-lowered-views ((xv image))
				 (unfold xv (device-vec-ptr (tensor-vec dense)) (tensor-initial-offset dense) g)))
			     dense))
		   (push i2 out))
		 (push i0 out)))
	finally (return (nreverse out))))

(defmethod on-finalizing-compiling ((device-name (eql 'CUDATensor)) iseq-fw iseq-bw)
  (if *cuda-available-p*
      (flet ((rewrite (iseq)
	       (multiple-value-bind (iseq sums)
		   (fuse-zero-fill+sum (fuse-matmul+bias (fuse-move+arith (fuse-unfold-chain iseq))))
		 (fold-grad-adder (fuse-gemm+relu (fuse-relu (fuse-softmax-chain iseq sums)))))))
	(let ((fw (rewrite iseq-fw))
	      (bw (and iseq-bw (rewrite iseq-bw))))
	  (alias-immutable-sv4bw fw bw)))
      (values iseq-fw iseq-bw)))

;; ---- LayerNorm over the last axis as ONE launch (csrc/softmax.cu `layernorm_kernel`).  The GPT-2 example
;; spells it as eleven nodes (`LayerNorm-Revisit`, examples/gpt-2/utils.lisp:114-127: two !mean, !sub, !expt,
;; !add eps, !sqrt, !div, !mul g, !add b); the kernel evaluates the same operations in the same order with
;; correctly rounded div/sqrt, so the result agrees with the node chain to the last few ulp.  Inference only
;; (the example never differentiates it): the backward is left to the node chain.
(define-op (CUDALayerNormNode (self eps)
	    :slots ((eps :initarg :eps :initform 1.0e-5 :reader eps-of))
	    :where (X[~ n] G[n] B[n] Out[~ n] -> Out[~ n])
	    :forward ((self x g b out)
		      (let* ((cols (car (last (shape x))))
			     (rows (/ (apply #'* (shape x)) cols)))
			(layernorm (device-vec-ptr (tensor-vec x)) (tensor-initial-offset x)
				   (device-vec-ptr (tensor-vec g)) (tensor-initial-offset g)
				   (device-vec-ptr (tensor-vec b)) (tensor-initial-offset b)
				   (device-vec-ptr (tensor-vec out)) (tensor-initial-offset out)
				   rows cols (coerce (eps-of self) 'single-float))
			out))
	    :backward ((self dout)
		       (declare (ignore dout))
		       (error "CUDALayerNormNode: forward only; differentiate LayerNorm-Revisit instead")))
  (setf (slot-value self 'eps) eps))

(defun cuda-layer-norm (x g b &key (eps 1.0e-5) feeds-matmul)
  "Drop-in for `LayerNorm-Revisit` on dense CUDATensors whose last axis is a multiple of 4 and <= 1024
(the library answers UNSUPPORTED otherwise, which `defwf2` signals as `cuda-backend-error`).
FEEDS-MATMUL: the result is the untransposed left operand of the `!matmul` issued next (LinearLayer after the norm): the kernel
also writes that product's tensor-core operand plane (wf2cu_emit_operand_planes, one shot; same bits as the product's own
pre-pass; ignored when the very next launch is not that product)."
  (when feeds-matmul (emit-operand-planes 1))
  (call (CUDALayerNormNode eps) (->contiguous x) g b (make-input (shape x) nil :dtype (dtype x) :order (order x))))

;; ---- `!softmax` (`nn/activation.lisp:126-159`: five nodes and two disguised reductions) as ONE launch, optionally with the
;; `(!div w n)` that precedes it in the GPT-2 example folded in (`examples/gpt-2/utils.lisp:236-242`: PRE-DIVISOR = n; 0 =
;; none).  AXIS counts from the front like `!softmax :axis`.  Forward only, like CUDALayerNormNode.
(define-op (CUDASoftmaxNode (self axis avoid-overflow pre-divisor)
	    :slots ((axis :initform 1 :reader axis-of)
		    (avoid-overflow :initform t :reader avoid-overflow-of)
		    (pre-divisor :initform 0.0 :reader pre-divisor-of))
	    :where (X[~] Out[~] -> Out[~])
	    :forward ((self x out)
		      (with-lowered-views ((xv x) (ov out))
			(softmax-ex xv ov (axis-of self) (if (avoid-overflow-of self) 1 0)
				    (coerce (pre-divisor-of self) 'single-float)))
		      out)
	    :backward ((self dout)
		       (declare (ignore dout))
		       (error "CUDASoftmaxNode: forward only; differentiate !softmax instead")))
  (setf (slot-value self 'axis) axis
	(slot-value self 'avoid-overflow) avoid-overflow
	(slot-value self 'pre-divisor) pre-divisor))

(defun cuda-softmax (x &key (axis 1) (avoid-overflow t) (pre-divisor 0.0))
  "Drop-in for `(!softmax x :axis axis :avoid-overflow avoid-overflow)`; with PRE-DIVISOR n for `(!softmax (!div x n) ...)`."
  (call (CUDASoftmaxNode axis avoid-overflow pre-divisor) x (make-input (shape x) nil :dtype (dtype x) :order (order x))))

;; ---- the GPT-2 example's SelfAttention score block (`examples/gpt-2/utils.lisp:204-244`, PAST = nil) as TWO launches:
;;   P  = softmax over the HEADS axis of (Q K^T / N)   (`!div` by the last axis' length, `!softmax :avoid-overflow nil` whose
;;        default :axis 1 is the heads axis of a (batch head N N) tensor; no mask)           wf2cu_attn_heads_softmax
;;   out = merge-heads (P V)                                                                 wf2cu_attn_pv
;; instead of 3 split copies, 2 batched matmuls, !div, the 6-instruction softmax chain and the merge copy.  X is the
;; (batch N 3*n-head*head-dim) result of the c_attn projection.  The library takes n-head 12, head-dim 64, N a multiple of 128
;; (the example's 124M configuration) and answers UNSUPPORTED otherwise, which `defwf2` signals as `cuda-backend-error`: call
;; the example's own SelfAttention then.  Forward only, like CUDALayerNormNode.
;; KEEP-PROBABILITIES nil (the default): one call, wf2cu_attn_fused — the weights only ever exist un-normalised in a workspace of the
;; library (one pass over the heads, half the Q K^T work), the result is bit-identical.  t: the two launches above, with the
;; (batch head N N) weights kept in the node's PROBABILITIES slot for inspection.
(define-op (CUDASelfAttentionNode (self n-head &optional keep-probabilities)
	    :slots ((n-head :initform 12 :reader n-head-of)
		    (keep-probabilities :initform nil :reader keep-probabilities-of)
		    (probabilities :initform nil :accessor probabilities-of))
	    :where (X[batch n three-e] Out[batch n e] -> Out[batch n e])
	    :forward ((self x out)
		      (let* ((batch (observe (first (shape x))))
			     (n     (observe (second (shape x))))
			     (e     (observe (third (shape out))))
			     (heads (n-head-of self))
			     (p     (probabilities-of self)))
			(cond
			  ((not (keep-probabilities-of self))
			   (attn-fused (device-vec-ptr (tensor-vec x)) (tensor-initial-offset x) batch n heads (/ e heads)
				       (coerce n 'single-float)
				       (device-vec-ptr (tensor-vec out)) (tensor-initial-offset out)))
			  (t
			   (unless (and p (equal (shape p) (list batch heads n n)))
			     (setf p (make-tensor (list batch heads n n) :dtype (dtype x) :order (order x))
				   (probabilities-of self) p))
			   (attn-heads-softmax (device-vec-ptr (tensor-vec x)) (tensor-initial-offset x) batch n heads (/ e heads)
					       (coerce n 'single-float)
					       (device-vec-ptr (tensor-vec p)) (tensor-initial-offset p))
			   (attn-pv (device-vec-ptr (tensor-vec p)) (tensor-initial-offset p)
				    (device-vec-ptr (tensor-vec x)) (tensor-initial-offset x) batch n heads (/ e heads)
				    (device-vec-ptr (tensor-vec out)) (tensor-initial-offset out))))
			out))
	    :backward ((self dout)
		       (declare (ignore dout))
		       (error "CUDASelfAttentionNode: forward only; differentiate SelfAttention instead")))
  (setf (slot-value self 'n-head) n-head
	(slot-value self 'keep-probabilities) keep-probabilities))

(defun cuda-self-attention (x &key (n-head 12) feeds-matmul keep-probabilities)
  "Drop-in for the first value of `(SelfAttention x nil)` on a dense (batch N 3E) CUDATensor.
FEEDS-MATMUL: as in `cuda-layer-norm` (the output projection follows; the request reaches the second of the two launches
because the first one is not a producer and leaves the one-shot flag alone)."
  (when feeds-matmul (emit-operand-planes 1))
  (let ((e (/ (car (last (shape x))) 3)))
    (call (CUDASelfAttentionNode n-head keep-probabilities)
	  (->contiguous x)
	  (make-input (list (first (shape x)) (second (shape x)) e) nil :dtype (dtype x) :order (order x)))))

;; ---- independent instructions on a side stream: (with-cuda-branch (1) <launches>) ... (branch-join 1)
(defmacro with-cuda-branch ((b) &body body)
  "BODY's launches go to side stream B (1..3), ordered after everything issued so far; the main stream
waits for them at `(branch-join B)`.  Inside `with-cuda-graph` these become fork/join edges."
  `(progn (branch-begin ,b)
	  (unwind-protect (progn ,@body)
	    (branch-end ,b))))

;; ---- CUDA-graph replay (SURVEY §8f rank 1): after `set-adjustable-symbols` the iseq is a fixed
;; launch sequence; capture `(forward model)` once and replay it.
(defmacro with-cuda-graph ((exec-var) capture-form &body body)
  `(let ((,exec-var (with-foreign-object (out :pointer)
		      (graph-begin)
		      ,capture-form
		      (graph-end out)
		      (mem-ref out :pointer))))
     (unwind-protect (progn ,@body)
       (graph-destroy ,exec-var))))

;; ---- data-parallel MLP step (SURVEY §8e): one process per GPU; gradients live in one arena
;; (allocate parameters' grads as views of a single CUDATensor) so the bucket is one range.
(defun cuda-comm-init (rank world id-bytes)
  "ID-BYTES: the 128-byte NCCL unique id made by rank 0 (wf2cu_comm_unique_id) and shared by file/env."
  (with-foreign-object (id :uint8 128)
    (dotimes (i 128) (setf (mem-aref id :uint8 i) (aref id-bytes i)))
    (comm-init rank world id)))

(defun cuda-allreduce-gradients (grad-arena n-floats)
  "Sum the gradient arena over ranks on the library stream; the 1/world factor is applied by the
optimizer kernel (`grad_scale` of wf2cu_adam_step)."
  (allreduce-sum-f32 (device-vec-ptr (tensor-vec grad-arena)) (tensor-initial-offset grad-arena) n-floats))

;; ---- exchange + Adam as ONE kernel over NVLink peer memory (csrc/p2p.cu): reduce-scatter, Adam on this
;; rank's shard, parameter all-gather.  HANDLES-EXCHANGE is any host transport that all-gathers 192 bytes per
;; rank (file, socket, MPI): it receives this rank's (unsigned-byte 8) vector and returns a list of WORLD vectors.
(defun cuda-p2p-attach (rank world grad-arena param-arena handles-exchange)
  (with-foreign-objects ((mine :uint8 192) (grads :uint8 (* 64 8)) (params :uint8 (* 64 8)) (sigs :uint8 (* 64 8)))
    (p2p-export (device-vec-ptr (tensor-vec grad-arena)) mine)
    (p2p-export (device-vec-ptr (tensor-vec param-arena)) (inc-pointer mine 64))
    (p2p-signal-export (inc-pointer mine 128))
    (let* ((blob (make-array 192 :element-type '(unsigned-byte 8)))
	   (all  (progn (dotimes (i 192) (setf (aref blob i) (mem-aref mine :uint8 i)))
			(funcall handles-exchange blob))))
      (loop for r from 0 below world
	    for h in all
	    do (dotimes (i 64)
		 (setf (mem-aref grads  :uint8 (+ (* 64 r) i)) (aref h i)
		       (mem-aref params :uint8 (+ (* 64 r) i)) (aref h (+ 64 i))
		       (mem-aref sigs   :uint8 (+ (* 64 r) i)) (aref h (+ 128 i)))))
      (p2p-attach rank world (device-vec-ptr (tensor-vec grad-arena)) (device-vec-ptr (tensor-vec param-arena))
		  grads params sigs))))

(defun cuda-p2p-adam-step (param-arena m-arena v-arena n-floats lr-t-device &key (beta1 0.9) (beta2 0.999) (eps 1e-7))
  "After `backward` on every rank: the whole data-parallel update in one launch."
  (p2p-allreduce-adam-step (device-vec-ptr (tensor-vec param-arena)) 0 0
			   (device-vec-ptr (tensor-vec m-arena)) 0 (device-vec-ptr (tensor-vec v-arena)) 0
			   n-floats 0.0 (device-vec-ptr (tensor-vec lr-t-device)) beta1 beta2 eps))
