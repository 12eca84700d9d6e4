ssh test@127.0.0.1 "kill -9 3135"
ssh test@127.0.0.1 "kill -9 3136"
ssh test@127.0.0.1 "kill -9 3137"
kill -9 2946
