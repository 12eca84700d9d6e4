ssh test@127.0.0.1 "kill -9 5490"
ssh test@127.0.0.1 "kill -9 5491"
ssh test@127.0.0.1 "kill -9 5492"
ssh test@127.0.0.1 "kill -9 5493"
kill -9 5044
