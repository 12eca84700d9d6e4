ssh test@127.0.0.1 "kill -9 3628"
ssh test@127.0.0.1 "kill -9 3629"
ssh test@127.0.0.1 "kill -9 3630"
kill -9 2946
